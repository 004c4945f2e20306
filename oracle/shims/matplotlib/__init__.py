"""ORACLE shim: the reference imports matplotlib.pyplot at module import but never uses it on the hot path."""
