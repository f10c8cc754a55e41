"""evox_myc_b200 -- B200-native non_dominated_sort / crowding_distance (EvoX NSGA-II/III tell()).

Only the environmental-selection hot path of the reference is implemented here;
it sits behind the reference's own operator signatures.  See DESIGN.md and include/evoxb.h.
"""

from . import _lib  # noqa: F401
from .operators import (  # noqa: F401
    NonDominate,
    crowding_distance,
    crowding_distance_sort,
    non_dominate,
    non_dominated_sort,
    nsga2_select,
    nsga3_last_rank,
)
from .nsga3 import UniformSampling, nsga3_select  # noqa: F401
from .utils import dominate_relation  # noqa: F401

__all__ = [
    "non_dominated_sort",
    "crowding_distance",
    "crowding_distance_sort",
    "non_dominate",
    "NonDominate",
    "nsga2_select",
    "nsga3_last_rank",
    "nsga3_select",
    "UniformSampling",
    "dominate_relation",
]
