"""One namespace for the whole Inception family: ``inception_family.inception_v4``,
``inception_family.inception_v3_arg_scope`` ... — the role of the zoo's ``vgg/nets/inception.py:21-37``, which exists
only so that callers need a single import.  The networks and scopes live in ``inception_zoo`` (table-driven v1 - v4 and
Inception-ResNet-v2); ``inference`` / ``loss`` are the training wrapper of the Inception-v3 the reference actually
trains (``inception/inception_model.py``)."""
from . import inception_v3 as _trained
from . import inception_zoo as _zoo

inference, loss = _trained.inference, _trained.loss

_VERSIONS = ("v1", "v2", "v3", "v4", "resnet_v2")
__all__ = ["inference", "loss", "inception_arg_scope"]
inception_arg_scope = _zoo.inception_arg_scope
for _v in _VERSIONS:
    for _suffix in ("", "_base", "_arg_scope"):
        _name = f"inception_{_v}{_suffix}"
        # v1 / v2 share v3's scope, as the zoo's factory wires them (nets_factory.py:72-75)
        _obj = getattr(_zoo, _name, None) or (_zoo.inception_v3_arg_scope if _suffix == "_arg_scope" else None)
        if _obj is not None:
            globals()[_name] = _obj
            __all__.append(_name)
del _v, _suffix, _name, _obj
