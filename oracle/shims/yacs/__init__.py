"""ORACLE shim: minimal yacs.config.CfgNode (attribute dict, clone, merge_from_file, freeze)."""
