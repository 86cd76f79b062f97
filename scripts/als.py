"""`scripts.als` — same module name as the reference; exports the B200 ALS."""
import b200als as _impl

ALS = _impl.ALS
SCALE_FACTOR = _impl.SCALE_FACTOR
EPS = _impl.EPS
__all__ = ["ALS", "SCALE_FACTOR", "EPS"]
