# -*- coding: utf-8 -*-
"""
Host utilities mirrored from the reference's acoss/utils.py: `log` (:24-35) and
`create_dataset_filepaths` (:87-102).  Same names, arguments and error behaviour.

The free functions of the reference's `acoss/algorithms/utils/` package (get_csm, csm_to_binary, getW, ...) are re-exported
here from `algorithm_utils.py`; they run on the device through the C-ABI.
"""
import logging

import pandas as pd

from . import algorithm_utils as _au
from .algorithm_utils import *          # noqa: F401,F403  (the reference's free-function names)

__all__ = ["log", "create_dataset_filepaths"] + list(_au.__all__)


def log(log_file):
    """Returns a logger object with predefined settings (acoss/utils.py:24-35)."""
    root_logger = logging.getLogger(__name__)
    root_logger.setLevel(logging.DEBUG)
    if not root_logger.handlers:
        fmt = logging.Formatter('%(asctime)s - %(name)s - %(levelname)s - %(message)s')
        fh = logging.FileHandler(log_file)
        fh.setFormatter(fmt)
        root_logger.addHandler(fh)
        ch = logging.StreamHandler()
        ch.setFormatter(fmt)
        root_logger.addHandler(ch)
    return root_logger


def create_dataset_filepaths(dataset_csv, root_audio_dir, file_format=".mp3"):
    """`root_audio_dir/work_id/track_id<file_format>` per csv row (acoss/utils.py:87-102).
    `root_audio_dir` is string-concatenated: it needs its trailing '/' (SURVEY Q14).
    Raises IOError for columns other than work_id / track_id."""
    dataset = pd.read_csv(dataset_csv, dtype=str)
    warn = ("Wrong input dataset csv annotation file '%s'. Expected a csv file with the columns of key "
            "'work_id', 'track_id'" % dataset_csv)
    for key in dataset.keys().tolist():
        if key not in ['work_id', 'track_id']:
            raise IOError(warn)
    return [root_audio_dir + str(w) + "/" + str(t) + file_format for w, t in zip(dataset.work_id, dataset.track_id)]
