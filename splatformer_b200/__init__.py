"""B200-native differentiable Gaussian-splat render (drop-in for the gsplat v0.1.11 path of SplatFormer)."""
__version__ = "0.1.0"
