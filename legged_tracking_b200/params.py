"""Minimal class-attribute config trees with the surface the reference uses from `params_proto` (pinned 2.10.5 in the
reference's setup.py:14, not vendored there): `class X(PrefixProto, cli=False)` with nested classes, configuration by
plain attribute assignment on the classes, and `vars(X)` returning a dict copy of the public attributes
(legged_robot_trajectory_tracking.py:1863-1865 relies on that)."""


_RAW_DICT = type.__dict__["__dict__"]


class Meta(type):
    def __new__(mcs, name, bases, ns, **kwargs):
        return super().__new__(mcs, name, bases, ns)

    def __init__(cls, name, bases, ns, **kwargs):
        super().__init__(name, bases, ns)

    @property
    def __dict__(cls):            # vars(Cls) -> public attributes only (what params_proto exposes)
        out = {}
        for klass in reversed(cls.__mro__):
            if klass is object:
                continue
            for k, v in _RAW_DICT.__get__(klass).items():
                if k.startswith("_") or isinstance(v, (classmethod, staticmethod, property)):
                    continue
                if callable(v) and not isinstance(v, type):
                    continue
                out[k] = v
        return out


class PrefixProto(metaclass=Meta):
    pass


class ParamsProto(metaclass=Meta):
    pass


def to_dict(cls):
    """Recursive plain-dict view of a config class tree (used for parameters.pkl-style dumps)."""
    out = {}
    for k, v in vars(cls).items():
        out[k] = to_dict(v) if isinstance(v, type) else v
    return out
