"""layerwise_b200 -- Blackwell-native compressed-gradient data-parallel training.

Capabilities follow the AAAI'20 layer-wise compression study code
(aritra-dutta/layer-wise-aaai20): the CIFAR-10 ``dawn.py`` recipe (ResNet-9,
AlexNet, VGG-16), the ImageNet ResNet-50 trainer, six gradient compressors
(Top-k, Random-k, Threshold-v, adaptive threshold, TernGrad, QSGD / random
dithering) applied layer-wise or to the entire model, error feedback, a
Random-k sparsifying DDP wrapper and a dense DDP wrapper.

Layout::

    ops/       compressors (CPU oracle + fused CUDA), SGD, BN, GEMM/conv, loss
    parallel/  process groups, peer-memory runtime, bucket planner, exchange
    models/    graph DSL, ResNet-9, AlexNet, VGG, ResNet-18..152
    data/      CIFAR / ImageNet pipelines (+ synthetic)
    train/     CIFAR (dawn) and ImageNet training loops
    utils/     timers, loggers, meters, schedules, checkpoints, fp16/bf16 masters
    csrc/      sm_100a CUDA + C++ sources of the native library
"""
__version__ = "0.1.0"
