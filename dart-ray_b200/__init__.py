"""dart-ray_b200 — B200-native engine for the DART-Ray ray-tracing hot path."""
__version__ = "0.1.0"
