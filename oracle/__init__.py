"""ORACLE — TEST INFRASTRUCTURE ONLY (**parity unpinned**; see pcc_oracle.cpp header).

CPU restatement of the reference's encode/decode path.  Importable only from tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
"""
from . import prims, entropy, codec  # noqa: F401
from .codec import OracleCodec  # noqa: F401
