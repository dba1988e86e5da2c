"""`from clusten import CLUSTENQKFunction, CLUSTENAVFunction, CLUSTENWFFunction` — the import line the reference's
models use (reference clusten/__init__.py:6; call sites models/aff_transformer.py:11, models/mixres_neighbour.py:10).
Put this directory's parent (`ml-autofocusformermod_b200/`) on sys.path in place of the reference's repo root."""
from affb200.clusten import CLUSTENQKFunction, CLUSTENAVFunction, CLUSTENWFFunction  # noqa: F401
