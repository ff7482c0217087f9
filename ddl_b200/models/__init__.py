from .spec import ModelSpec, VarSpec, make_spec, REFERENCE_SPEC  # noqa: F401
