"""sequenza_b200: the cellularity x ploidy likelihood path of sequenza 3.0.0
(sequenza.fit -> baf.model.fit -> baf.bayes, and mufreq.bayes) on B200 GPUs.

The Python functions mirror the reference's R interface for this path and call
hand-written sm_100a kernels through the C ABI in include/sequenza_b200.h.  There is
no CPU implementation: without libsequenza_b200.so or without a GPU every compute call
raises."""
from ._lib import SQZ_MODE_FAITHFUL, SQZ_MODE_FAST, SO_PATH, lib  # noqa: F401
from .bayes import (baf_bayes, baf_grid_sums, baf_model_fit, baf_model_fit_batch,  # noqa: F401
                    finish_model_fit, mufreq_bayes, mufreq_grid_sums, mufreq_model_fit,
                    r_seq)
from .fit import sequenza_fit  # noqa: F401
from .model import (baf_model_points, baf_types_matrix, device_model_points,  # noqa: F401
                    mufreq_model_points, mufreq_types_matrix, theoretical_baf,
                    theoretical_depth_ratio, theoretical_mufreq)
from .segments import segment_breaks, segment_stats  # noqa: F401
from .solutions import alternative_cp_solutions, get_ci  # noqa: F401

__version__ = "0.1.0"
