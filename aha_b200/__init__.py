"""aha_b200 -- B200-native drop-in for the Aha::AC scan path of chenkovsky/aha."""
from ._lib import AhaError, build, lib  # noqa: F401
from .ac import (AC, COUNT_ONLY, HIT_DTYPE, IGNORE_CASE, INTERSECTABLE, LONGEST, SEP, SEP_CODEPOINT, TRIE_LONGEST, UNITS_CHAR,  # noqa: F401
                 Cedar, ScanResult, sep_bitmap)
