"""B200-native ORB front-end: drop-in for ORB-SLAM3's ORBextractor::operator() and the
ORBmatcher descriptor-distance search (reference: zhoujoey/ORB-SLAM3-Mac).

Product code is C++/CUDA behind the C ABI in include/orbx.h (built into
orb-slam3-mac_b200/liborbx_b200.so).  This Python package is only the thin host-side mirror
used by tests/ and bench.py: ctypes bindings plus classes with the reference's names.
There is no CPU fallback: importing the bindings without the built library raises.
"""
from . import synth  # noqa: F401
