"""Polling evaluators: precision@1 (CIFAR-10) / precision@1 + recall@5 (ImageNet) on the newest checkpoint."""
