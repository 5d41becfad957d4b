"""egoscaler_b200 -- B200-native (sm_100a) implementation of EgoScaler's trajectory-generation hot path."""
__version__ = "0.1.0"
