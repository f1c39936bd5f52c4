"""Training loops: CIFAR-10 (dawn recipe) and ImageNet (phased ResNet-50 recipe)."""
