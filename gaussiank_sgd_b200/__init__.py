"""B200-native sparsified data-parallel SGD with the capabilities of sb17v/GaussianK-SGD."""
__version__ = "0.1.0"
