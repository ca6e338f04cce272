from .core import (byzantine_tolerance, fedavg, fltracer_anomalies, fltrust, gmm_filter,  # noqa: F401
                   krum_select, median, pairwise_sqdist, scion_dequant_scores, scion_quantize, scion_scores_torch, scionfl, scionfl_keep, shieldfl,
                   shieldfl_weights, trimmed_mean, weighted_sum)
from .hyper import hyper_round_torch, hyper_step_torch  # noqa: F401
