"""Input pipelines: CIFAR-10 binary reader + augmentation, ImageNet TFRecord reader + distortion,
synthetic generators of both shapes, and the offline dataset builders."""
