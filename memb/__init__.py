"""Drop-in alias so that `import memb` resolves to the B200-native package
(same names as python/memb/__init__.py:1-4 of the reference)."""
from memb_b200 import Builder, Reader, ReadersUnion, available_compression_strategies  # noqa: F401
