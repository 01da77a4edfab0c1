"""The six trained networks, the nets-factory registry and the vendored model zoo (Inception family, slim ResNets,
MobileNets, NASNet / PNASNet, GAN nets, LeNet, OverFeat)."""
