# -*- coding: utf-8 -*-
"""
acoss_b200 -- B200-native implementation of acoss's all-pairs cover-song-similarity hot path
behind the reference's CoverAlgorithm API.  The directory is named `acoss-1_b200`; import it as
`acoss_b200` through the shim module at the repository root.
"""
from . import coverid, synth, featio, utils, algorithm_utils
from ._lib import AcossError, Context, load_library, LIB_PATH
from .algorithm_template import CoverAlgorithm
from .coverid import algorithm_names, benchmark
from .latefusion_chen import ChenFusion
from .rqa_serra09 import Serra09
from .simple_silva import Simple
from .ftm2d import FTM2D
from .earlyfusion_traile import EarlyFusion

__version__ = "0.1.0"
