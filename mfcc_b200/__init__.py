"""B200-native MFCC extractor behind the stage API of lambdaconcept/mfcc."""
from .config import MFCCConfig, MODE_INT, MODE_FP32  # noqa: F401
