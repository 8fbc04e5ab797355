"""mgl_b200 -- B200-native Smith-Waterman affine-gap aligner, a drop-in for microsoft/mgl's mgl_sw path."""
from .smithwaterman import (BatchResult, MicrosoftSmithWaterman, SWNativeAlignerResult, SWOverhangStrategy,  # noqa: F401
                            SWParameters, align_one_text)
