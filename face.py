"""Drop-in module name of the reference (`from face import FaceDet`)."""
from depth_estimation_b200.face import FaceDet  # noqa: F401
