"""Data-parallel machinery: process groups, peer memory, bucket planning, gradient exchange."""
from . import comm  # noqa: F401
from .exchange import (CompressionConfig, GradientExchanger, layerwise_compressed_comm,  # noqa: F401
                       entiremodel_compressed_comm, all_reduce, canonical_compress)
