"""Drop-in mirror of the reference's ``ReSIDE`` package for the depth-estimation forward path.

Same module / class / parameter names as /root/reference/ReSIDE, so the
reference's ``.pth`` checkpoints load unchanged; arithmetic runs in libede_sm100.so.
"""
