"""Minimal stand-in for `params_proto==2.10.5` (pinned in the reference's setup.py:14, not vendored).

Oracle shim, test infrastructure only.  The reference uses it purely as a class-attribute
namespace: `class Cfg(PrefixProto, cli=False)` trees mutated imperatively, and
`vars(Cfg.reward_scales)` to obtain a *dict copy* of the public attributes
(legged_robot_trajectory_tracking.py:1863-1865).
"""


class Meta(type):
    def __new__(mcs, name, bases, ns, **kwargs):
        return super().__new__(mcs, name, bases, ns)

    def __init__(cls, name, bases, ns, **kwargs):
        super().__init__(name, bases, ns)

    @property
    def __dict__(cls):
        out = {}
        for klass in reversed(cls.__mro__):
            if klass in (object, ):
                continue
            for k, v in type.__dict__["__dict__"].__get__(klass).items():
                if not k.startswith("_"):
                    out[k] = v
        return out

    def __vars__(cls):
        return cls.__dict__


class ParamsProto(metaclass=Meta):
    pass


class PrefixProto(ParamsProto):
    pass


class Proto:
    def __init__(self, default=None, **kw):
        self.default = default


class Flag(Proto):
    pass
