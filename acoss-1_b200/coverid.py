# -*- coding: utf-8 -*-
"""
Interface to run the cover id algorithms for benchmarking: mirror of the reference's
acoss/coverid.py (`benchmark` :22-142, `algorithm_names` :19).
Repairs: Q1 -- both "EarlyFusionTraile" (the listed name) and "EarlyFusion"
(the name the reference's dispatch tests for) run the EarlyFusion algorithm.
"""
import time

from .utils import log

__all__ = ['benchmark', 'algorithm_names']

_LOG_FILE_PATH = "acoss.coverid.log"
_LOGGER = None

# list the available cover song identification algorithms in acoss
algorithm_names = ["Serra09", "EarlyFusionTraile", "LateFusionChen", "FTM2D", "SiMPle"]


def _logger():
    global _LOGGER
    if _LOGGER is None:
        _LOGGER = log(_LOG_FILE_PATH)
    return _LOGGER


def benchmark(dataset_csv, feature_dir, feature_type="hpcp", algorithm="Serra09", shortname="covers80",
              parallel=True, n_workers=-1, **algo_kwargs):
    """Benchmark a specific cover id algorithm with a given input dataset annotation csv file.
    Same arguments as the reference; `parallel` / `n_workers` are accepted and ignored (the GPU
    batches the pair grid).  Extra keyword arguments go to the algorithm constructor.
    Returns {similarity_type: (MR, MRR, MDR, MAP, tops)} (the reference returns None).

    Raises NotImplementedError when the algorithm name is unknown (coverid.py:46-51)."""
    if algorithm not in algorithm_names + ["EarlyFusion"]:
        warn = ("acoss.coverid: Couldn't find '%s' algorithm in acoss \
                                Available cover id algorithms are %s " % (algorithm, str(algorithm_names)))
        _logger().debug(warn)
        raise NotImplementedError(warn)
    _logger().info("Running acoss cover identification benchmarking for the algorithm - '%s'" % algorithm)
    start_time = time.monotonic()
    results = {}
    if algorithm == "Serra09":
        from .rqa_serra09 import Serra09
        algo = Serra09(dataset_csv=dataset_csv, datapath=feature_dir, chroma_type=feature_type,
                       shortname=shortname, **algo_kwargs)
        _logger().info('Computing pairwise similarity...')
        algo.all_pairwise(parallel, n_cores=n_workers, symmetric=True)
        algo.normalize_by_length()
    elif algorithm == "LateFusionChen":
        from .latefusion_chen import ChenFusion
        algo = ChenFusion(dataset_csv=dataset_csv, datapath=feature_dir, chroma_type=feature_type,
                          shortname=shortname, **algo_kwargs)
        _logger().info('Computing pairwise similarity...')
        algo.all_pairwise(parallel, n_cores=n_workers, symmetric=True)
        algo.normalize_by_length()
        algo.do_late_fusion()
    elif algorithm in ("EarlyFusion", "EarlyFusionTraile"):
        from .earlyfusion_traile import EarlyFusion
        algo = EarlyFusion(dataset_csv=dataset_csv, datapath=feature_dir, chroma_type=feature_type,
                           shortname=shortname, **algo_kwargs)
        algo.preload_features()                       # all tracks in one device batch (the reference loops over load_features)
        for i in range(len(algo.filepaths)):
            algo.load_features(i)
        _logger().info('Feature loading done...')
        algo.all_pairwise(parallel, n_cores=n_workers, symmetric=True)
        algo.do_late_fusion()
    elif algorithm == "FTM2D":
        from .ftm2d import FTM2D
        algo = FTM2D(dataset_csv=dataset_csv, datapath=feature_dir, chroma_type=feature_type,
                     shortname=shortname, **algo_kwargs)
        algo.preload()                                # all shingles in one device batch
        for i in range(len(algo.filepaths)):
            algo.load_features(i)
        _logger().info('Feature loading done...')
        algo.all_pairwise(parallel, n_cores=n_workers, symmetric=True)
    elif algorithm == "SiMPle":
        from .simple_silva import Simple
        algo = Simple(dataset_csv=dataset_csv, datapath=feature_dir, chroma_type=feature_type,
                      shortname=shortname, **algo_kwargs)
        algo.preload()                                # WIN/SKIP means, smoothing, normalisation of all tracks in one device batch
        for i in range(len(algo.filepaths)):
            algo.load_features(i)
        _logger().info('Feature loading done...')
        algo.all_pairwise(parallel, n_cores=n_workers, symmetric=False)
    _logger().info('Running benchmark evaluations on the given dataset - %s' % dataset_csv)
    for similarity_type in list(algo.Ds.keys()):
        results[similarity_type] = algo.getEvalStatistics(similarity_type)
    algo.cleanup_memmap()
    _logger().info("acoss.coverid benchmarking finsihed in %s" % (time.monotonic() - start_time))
    return results
