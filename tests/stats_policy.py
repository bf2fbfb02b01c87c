"""The reference's policy for statistical tests (test/stats.py), restated for pytest.

  * a statistical test is a function of a `seed` that returns a TestResult (p-value + report)   stats.py:84-92
  * it is run once; p > 0.05 passes at once                                                      stats.py:101-110
  * a p-value that no four further perfect runs could rescue fails at once                       stats.py:111-113
  * otherwise it is run four more times with fresh seeds and the five p-values are combined
    with Fisher's method                                                                         stats.py:94-99, 114-126
  * the (combined) p-value must exceed the global reporting threshold, 1e-5                      stats.py:128-131;
                                                                                                 test/config.py
The report_known_* functions are the reference's reportKnown* (stats.py:151-268): KS against a known CDF /
Gaussian, chi-square against known rates, two-sample KS and chi-square.
"""
import functools
import math
import random

import numpy as np
import scipy.stats as stats

GLOBAL_REPORTING_THRESHOLD = 1e-5


class TestResult(object):
    __test__ = False          # not a pytest class

    def __init__(self, pval, report):
        self.pval, self.report = float(pval), report

    def __str__(self):
        return self.report


def fisher_method(pvals):
    if any(p == 0 for p in pvals):
        return 0.0
    chisq = -2.0 * sum(math.log(p) for p in pvals)
    return float(stats.chi2.sf(chisq, 2 * len(pvals)))


def repeat_test(func, seed, *args, **kwargs):
    rng = random.Random(seed)
    result = func(*args, seed=rng.randint(1, 2 ** 31 - 1), **kwargs)
    assert isinstance(result, TestResult)
    if result.pval > 0.05:
        return result
    if fisher_method([result.pval] + [1.0] * 4) < GLOBAL_REPORTING_THRESHOLD:
        return result                                  # hopeless
    results = [result] + [func(*args, seed=rng.randint(1, 2 ** 31 - 1), **kwargs) for _ in range(4)]
    pval = fisher_method([r.pval for r in results])
    report = "Test failing consistently\n" + "\n".join(r.report for r in results) + "\nOverall P value: %s" % (pval,)
    return TestResult(pval, report)


def report_test(result):
    assert result.pval > GLOBAL_REPORTING_THRESHOLD, str(result)


def statistical_test(f=None, seed=20260101):
    """Decorator: `f(..., seed)` returns a TestResult; the wrapped test applies the retry + Fisher policy."""
    def deco(fn):
        @functools.wraps(fn)
        def wrapped(*args, **kwargs):
            kwargs.pop("seed", None)
            report_test(repeat_test(fn, seed, *args, **kwargs))
        # pytest must not see the `seed` parameter as a fixture
        import inspect
        sig = inspect.signature(fn)
        wrapped.__signature__ = sig.replace(parameters=[p for n, p in sig.parameters.items() if n != "seed"])
        return wrapped
    return deco(f) if f is not None else deco


def _explain(observed):
    o = np.asarray(observed, dtype=np.float64)
    pct = np.percentile(o, [0, 5, 25, 50, 75, 95, 100]) if o.size else []
    return "Observed: %4d samples with mean %4.3f, stddev %4.3f, percentiles (0,5,25,50,75,95,100) %s" % (
        o.size, o.mean() if o.size else float("nan"), o.std() if o.size else float("nan"), np.round(pct, 3).tolist())


def report_known_continuous(expected_cdf, observed, descr=None):
    K, pval = stats.kstest(np.asarray(observed, dtype=np.float64), expected_cdf)
    return TestResult(pval, "\n".join(["Expected: %4d samples from %s" % (len(observed), descr), _explain(observed),
                                       "K stat  : %s" % (K,), "P value : %s" % (pval,)]))


def report_known_gaussian(mean, std, observed):
    return report_known_continuous(stats.norm(loc=mean, scale=std).cdf, observed, "N(%s,%s)" % (mean, std))


def report_same_continuous(observed1, observed2):
    D, pval = stats.ks_2samp(np.asarray(observed1, dtype=np.float64), np.asarray(observed2, dtype=np.float64))
    return TestResult(pval, "\n".join(["Expected samples from the same distribution", _explain(observed1), _explain(observed2),
                                       "D stat  : %s" % (D,), "P value : %s" % (pval,)]))


def report_known_discrete(expected_rates, observed):
    """expected_rates: [(item, rate)] (normalised here); observed: list of items"""
    items = [i for i, _ in expected_rates]
    rates = np.array([r for _, r in expected_rates], dtype=np.float64)
    obs = list(observed)
    counts = np.array([sum(1 for o in obs if o == i) for i in items], dtype=np.float64)
    stray = len(obs) - counts.sum()
    if stray > 0:
        return TestResult(0.0, "%d observations outside the support %r" % (stray, items))
    exp = rates / rates.sum() * len(obs)
    keep = exp > 0
    if np.any(counts[~keep] > 0):
        return TestResult(0.0, "observed an outcome of probability zero")
    chisq, pval = stats.chisquare(counts[keep], exp[keep])
    return TestResult(pval, "\n".join(["Expected: %s" % np.round(exp, 1).tolist(), "Observed: %s" % counts.astype(int).tolist(),
                                       "Chi^2   : %s" % (chisq,), "P value : %s" % (pval,)]))


def report_known_counts(expected_probs, counts):
    """chi-square of a vector of counts against known cell probabilities"""
    counts = np.asarray(counts, dtype=np.float64)
    p = np.asarray(expected_probs, dtype=np.float64)
    exp = p / p.sum() * counts.sum()
    keep = exp > 0
    if np.any(counts[~keep] > 0):
        return TestResult(0.0, "observed an outcome of probability zero")
    chisq, pval = stats.chisquare(counts[keep], exp[keep])
    return TestResult(pval, "Expected: %s\nObserved: %s\nChi^2   : %s\nP value : %s" % (np.round(exp, 1).tolist(),
                                                                                      counts.astype(int).tolist(), chisq, pval))


def report_same_discrete_counts(counts1, counts2):
    c = np.array([counts1, counts2], dtype=np.float64)
    c = c[:, c.sum(0) > 0]
    chisq, pval, _, _ = stats.chi2_contingency(c)
    return TestResult(pval, "Expected two samples from the same discrete distribution\nObserved: %s\nObserved: %s\nChi^2   : %s\nP value : %s"
                      % (c[0].astype(int).tolist(), c[1].astype(int).tolist(), chisq, pval))
