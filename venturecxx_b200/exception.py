"""VentureException, mirroring python/lib/exception.py:21-66 of the reference
(same constructor, attributes and JSON form) so callers can catch the same
type the reference raises on this path."""


class VentureException(Exception):
    def __init__(self, exception, message=None, **kwargs):
        if message is None:
            message = exception
            exception = ""
        Exception.__init__(self, exception, message)
        self.exception = exception
        self.message = message
        self.data = kwargs

    def to_json_object(self):
        d = {"exception": self.exception, "message": self.message}
        d.update(self.data)
        return d

    @classmethod
    def from_json_object(cls, json_object):
        data = dict(json_object)
        return cls(data.pop("exception"), data.pop("message"), **data)

    def __str__(self):
        return "*** %s: %s\n%s" % (self.exception, self.message, self.data)

    __repr__ = __str__


# infer/subsampled_mh.py:26-28 (python/lib/exception.py in the reference): what subsampled_mh_check_applicability warns with
class SubsampledScaffoldNotEffectiveWarning(UserWarning):
    pass


class SubsampledScaffoldNotApplicableWarning(UserWarning):
    pass


class SubsampledScaffoldStaleNodesWarning(UserWarning):
    pass
