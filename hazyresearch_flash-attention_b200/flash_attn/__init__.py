"""B200-native drop-in for the `flash_attn` package of HazyResearch flash-attention v0.1
(reference: flash_attn/__init__.py is empty; the public API lives in the submodules)."""
__version__ = "0.1.0+b200"
