"""normflows -- drop-in for the reference package of the same name
(normflows/normflows/ in benlevyx/vi-normflows), with the planar-flow hot path on
B200 CUDA kernels (libnfcore).  Same module names, same public functions, NumPy in /
NumPy out.  There is no CPU fallback: functions that run kernels raise
``nfcore.NFError`` when no CUDA device (or no built libnfcore.so) is present.
"""
__version__ = '0.0.1'      # normflows/setup.py:4-6
