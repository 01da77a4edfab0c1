"""NHWC functional ops with TensorFlow semantics (``functional``), autograd wrappers of the native sm_100a kernels
(``native_ops``, ``conv_native``, ``zoo_native``) and branch-level stream concurrency (``streams``)."""
