from .mnist import load_mnist_data, synthetic_mnist
from .synthetic import ClientShards, make_client_shards

__all__ = ["load_mnist_data", "synthetic_mnist", "ClientShards", "make_client_shards"]
