"""One namespace for the whole Inception family — the role of the reference's ``vgg/nets/inception.py:21-37``,
which re-exports every version and its arg scope so that callers write ``inception.inception_v4`` /
``inception.inception_v3_arg_scope``.  (``inference`` / ``loss`` are the training wrapper of the model the reference
actually trains, ``inception/inception_model.py``.)"""
from .inception_v3 import inference, loss  # noqa: F401
from .inception_zoo import (  # noqa: F401
    inception_arg_scope,
    inception_resnet_v2,
    inception_resnet_v2_arg_scope,
    inception_resnet_v2_base,
    inception_v1,
    inception_v1_base,
    inception_v2,
    inception_v2_base,
    inception_v3,
    inception_v3_arg_scope,
    inception_v3_base,
    inception_v4,
    inception_v4_arg_scope,
    inception_v4_base,
)

inception_v1_arg_scope = inception_v3_arg_scope
inception_v2_arg_scope = inception_v3_arg_scope
