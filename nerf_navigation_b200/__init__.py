"""nerf_navigation_b200 — B200-native (sm_100a) neural-field hot path behind the reference's operator API.

Sub-packages mirror the reference's top-level modules: gridencoder, raymarching, shencoder, ffmlp, activation,
encoding, nerf.  `install_as_reference_modules()` registers them under the reference's import names so that
`nav/` and `nerf/renderer.py` of stein7/nerf-navigation import them unchanged.
"""
__version__ = "0.1.0"
