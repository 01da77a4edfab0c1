"""`slim` namespace (role of the reference's ``inception/slim/slim.py:25-30``): layer library, variable
store, argument scopes and losses, importable as ``from distributed_tensorflow_models_b200 import slim``."""
from . import losses, ops, scopes, variables  # noqa: F401
from .scopes import add_arg_scope, arg_scope, has_arg_scope  # noqa: F401
from .variables import VariableDeviceChooser, VariableStore  # noqa: F401
