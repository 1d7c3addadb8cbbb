"""columba_b200: B200-native approximate-matching hot path of Columba Vanilla."""
from .api import (ClbError, Index, Mapper, BatchResult, build_index, load_library, SCHEMES_DIR, DEFAULT_LIB,  # noqa: F401
                  COUNTER_NAMES, STAT_NAMES, STAGE_NAMES, cigar_string)
