"""evox_myc_b200 -- B200-native non_dominated_sort / crowding_distance (EvoX NSGA-II/III tell()).

Only the environmental-selection hot path of the reference is implemented here;
it sits behind the reference's own operator signatures.  See DESIGN.md and include/evoxb.h.
"""

from . import _lib  # noqa: F401
from .operators import non_dominated_sort  # noqa: F401
from .utils import dominate_relation  # noqa: F401

__all__ = ["non_dominated_sort", "dominate_relation"]
