"""argv-compatible replacements of the two reference tool scripts (SURVEY §8b "CLI contract"): drop
`shims/inference.py` in place of SwinIR/inference.py and `shims/demo.py` in place of Restormer/demo.py and the
UNCHANGED `Tool._get_cmd` subprocess path (AgenticIR/executor/tool.py:72-78) reaches the B200 kernels."""
