"""Import-compatibility alias for code written against the reference package layout.

``import CommEfficient.fed_aggregator``, ``from CommEfficient.utils import parse_args``,
``from CommEfficient.data_utils import FedCIFAR10`` ... resolve to the corresponding ``commefficient_b200`` modules
(same public names: ``FedModel``, ``FedOptimizer``, ``get_server_update``, ``process_batch``, ``local_step``,
``forward_grad``, ``parse_args``, the datasets, the models).  Nothing is defined here.
"""
import sys as _sys

from commefficient_b200 import (csvec, cv_train, data_utils, fed_aggregator, fed_worker, gpt2_train, models,  # noqa: F401
                                utils)

for _name in ("csvec", "cv_train", "data_utils", "fed_aggregator", "fed_worker", "gpt2_train", "models", "utils"):
    _sys.modules[__name__ + "." + _name] = globals()[_name]
for _sub in ("fed_sampler", "fed_dataset", "transforms"):
    _mod = getattr(data_utils, _sub, None)
    if _mod is not None:
        _sys.modules[__name__ + ".data_utils." + _sub] = _mod
