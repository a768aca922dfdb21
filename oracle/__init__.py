"""oracle/ — TEST INFRASTRUCTURE (never imported by the product package).

`oracle.lib()`      -> numpy-friendly wrapper around oracle/_build/libppo_oracle.so (plain-C restatement)
`oracle.ref()`      -> wrapper around oracle/_ref/libppo_ref.so (the compiled reference), or None
                       when it has not been built (it can only be built where /root/reference exists).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use this.
"""
from .loader import lib, ref, build_oracle, build_ref, ORACLE_DIR  # noqa: F401
