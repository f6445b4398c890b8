"""Import shim used ONLY by the reference arm of bench.py.

The reference imports ``matplotlib.pyplot`` unconditionally at module import
(``ddlb/benchmark.py:13``) although ``run_benchmark`` never plots; matplotlib is not
installed in this image and there is no network. This stub satisfies the import
without modifying the reference. Any attempt to actually plot raises.
"""
__version__ = "0.0-shim"
