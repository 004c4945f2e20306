"""ORACLE shim: import-compatible stand-in for the un-vendored `pytorchvideo` package (see oracle/pv_restate.py)."""
__version__ = "0.1.5-restated"
