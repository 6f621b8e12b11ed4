"""Drop-in module name of the reference (`from depth_midas import DepthEstimator`)."""
from depth_estimation_b200.depth_midas import DepthEstimator, VidStream  # noqa: F401
