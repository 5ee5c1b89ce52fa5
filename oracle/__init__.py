# -*- coding: utf-8 -*-
"""
ORACLE - TEST INFRASTRUCTURE ONLY (CPU restatement of the reference's batched native path).

Importable from ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py`` (cpu_baseline /
``--impl reference``) only.  ``pyodesys_b200`` never imports this package.
"""
from .build import build_oracle, build_ref, OracleModule, RefAnyodeLU  # noqa: F401
