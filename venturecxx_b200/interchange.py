"""State interchange with the reference's Lite backend.

A device particle is written out as what `engine.trace.Trace.dump()` returns
(python/lib/engine/trace.py:164-166):

    (values, directives, foreign_sp_names)

* `directives`: {directive id: ["define", name, datum] | ["observe", datum, value]} with datums as stack
  dicts (python/lib/value/dicts.py) -- the VentureScript program the compiled model stands for
  (`Model.program`, models.py), macro-expanded the way the sivm hands it to a trace (lambda -> make_csp,
  if -> biplex; python/lib/sivm/macro.py:293-311);
* `values`: the value of every RANDOM application in the trace as a boxed stack dict, in the order
  `_dump_trace` leaves them (engine/trace.py:179-202, lite/serialize.py:20-83): directives are unevaluated
  last-to-first and every random node pushes its value, so the list is the REVERSE of evaluation order and
  `_restore_trace` pops from its end while it re-evaluates the directives first-to-last.  Observed nodes
  are random applications too: their (constrained) values are on the stack.

`Trace.restore(mk_trace, serialized, foreign_sps)` (engine/trace.py:168-171) then rebuilds a Lite trace that
holds exactly the particle's state -- the parity statement "identical traces" made literal, and the route
by which `engine.save/load`, `convert_model` and `to_lite()` (engine.py:377-418, trace_set.py:377-398) can
carry device particles.  `load` is the inverse: Lite dumps -> device particles.

Nothing here imports the reference; tests/test_interchange*.py restores the dumps with the real Lite.
"""
import numpy as np

from .exception import VentureException


# ---- stack dicts (python/lib/value/dicts.py) ------------------------------------------------
def symbol(s):
    return {"type": "symbol", "value": s}


def number(x):
    return {"type": "number", "value": float(x)}


def integer(x):
    return {"type": "integer", "value": int(x)}


def boolean(b):
    return {"type": "boolean", "value": bool(b)}


def quote(datum):
    return [symbol("quote"), datum]


def expression(e):
    """Nested python lists / str / numbers -> stack-dict expression, with the sivm's two macros."""
    if isinstance(e, str):
        return symbol(e)
    if isinstance(e, bool):
        return boolean(e)
    if isinstance(e, (int, float, np.integer, np.floating)):
        return number(e)
    if isinstance(e, dict):
        return e
    if not isinstance(e, (list, tuple)):
        raise VentureException("invalid_argument", "cannot turn %r into an expression" % (e,))
    e = list(e)
    if e and e[0] == "lambda":
        return [symbol("make_csp"), quote([expression(p) for p in e[1]]), quote(expression(e[2]))]
    if e and e[0] == "if":
        return [[symbol("biplex"), expression(e[1]), expression(["lambda", [], e[2]]),
                 expression(["lambda", [], e[3]])]]
    if e and e[0] == "quote":
        return quote(expression(e[1]))
    return [expression(x) for x in e]


def define(name, e):
    return ["define", name, expression(e)]


def observe(e, value):
    return ["observe", expression(e), value]


def as_directive_dict(program):
    """Directive ids count from 1 in program order (engine/trace.py keeps {did: directive})."""
    return dict((i + 1, d) for i, d in enumerate(program))


# ---- dump ------------------------------------------------------------------------------------
def dump_particle(ts, index, sites=None):
    """(values, directives, foreign_sp_names) of particle `index` of a B200TraceSet.
    `sites` may carry the already-read site matrix (n_sites x P) when many particles are dumped."""
    model = ts.model
    if not hasattr(model, "program"):
        raise VentureException("invalid_argument", "model %r has no VentureScript program" % (model,))
    if sites is None:
        sites = read_sites(ts)
    obs = ts.observations()
    program = model.program(obs)
    values = model.random_values(obs, sites[:, index])
    return (list(reversed(values)), as_directive_dict(program), set())


def dump(ts):
    """Every particle, as `TraceSet.dump` would hand them to engine.save (trace_set.py:377-385)."""
    sites = read_sites(ts)
    return [dump_particle(ts, i, sites) for i in range(ts.n_particles)]


def read_sites(ts):
    n = ts.num_sites()
    if n == 0:
        return np.zeros((0, ts.n_particles))
    return np.stack([ts.read_site(j) for j in range(n)])


# ---- load ------------------------------------------------------------------------------------
def load(ts, serialized):
    """Inverse of `dump`: one Lite dump per device particle.  The directives must be the program of the
    trace set's model on its current observations (the data are not re-read from the dump); latent values
    are written into the sites and the derived caches recomputed (vk_refresh)."""
    if len(serialized) != ts.n_particles:
        raise VentureException("invalid_argument", "need one dump per particle (%d), got %d"
                               % (ts.n_particles, len(serialized)))
    model = ts.model
    obs = ts.observations()
    program = as_directive_dict(model.program(obs))
    n = ts.num_sites()
    sites = np.empty((n, ts.n_particles))
    for i, (values, directives, _names) in enumerate(serialized):
        if _strip(directives) != _strip(program):
            raise VentureException("invalid_argument", "dump %d holds a different program than this model" % i)
        sites[:, i] = model.sites_from_random_values(obs, list(reversed(values)))
    for j in range(n):
        ts.write_site(j, sites[j])
    ts.refresh()


def _strip(directives):
    """Directive dicts compare by content (numbers as floats)."""
    def norm(x):
        if isinstance(x, dict):
            v = x.get("value")
            return (x.get("type"), float(v) if isinstance(v, (int, float, np.integer, np.floating)) and
                    not isinstance(v, bool) else norm(v) if isinstance(v, (list, tuple, dict)) else v)
        if isinstance(x, (list, tuple)):
            return tuple(norm(y) for y in x)
        return x
    return dict((int(k), norm(v)) for k, v in directives.items())


def unbox(v):
    return v["value"] if isinstance(v, dict) and "value" in v else v
