"""B200-native wavefront light-transport hot path behind the reference's Radeon Rays
IntersectionApi and path-tracer pass interface.

The directory name carries a hyphen (it mirrors the reference repo's name); import it through the
root-level loader module `mcrt_b200` (mcrt_b200.py), which registers this package under that name.
"""
__version__ = "0.1.0"
