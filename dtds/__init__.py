"""CLI-compatibility package: ``python -m dtds.distributed ...`` (reference entry point,
``Server/dtds/distributed.py:894-971``) resolves to ``fed_tgan_b200.parallel.launcher``."""
