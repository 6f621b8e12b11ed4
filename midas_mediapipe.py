"""Drop-in module name of the reference's stream script (midas_mediapipe.py:11-160): headless VidStream."""
from depth_estimation_b200.midas_mediapipe import VidStream  # noqa: F401
