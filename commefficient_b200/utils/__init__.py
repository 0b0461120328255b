"""``utils`` surface of the reference (/root/reference/CommEfficient/utils.py)
re-exported under the same names."""
from .args import (MODES, ERROR_TYPES, DATASET_CLASSES, parse_args, make_args,
                   validate_args, build_parser, num_classes_of_dataset)
from .logging import (Logger, TableLogger, TSVLogger, JSONLLogger, Timer,
                      DeviceTimer, PhaseTracer, make_logdir, union, profile_helper, sm2np, shms)
from .schedules import (PiecewiseLinear, Exp, steps_per_epoch, triangular_lr,
                        linear_decay_lr)
from .flat import (FlatParams, get_param_vec, set_param_vec, get_grad_vec,
                   get_grad, zero_grad, clip_grad)

fed_datasets = DATASET_CLASSES


def is_port_in_use(port: int) -> bool:
    """utils.py:46-49."""
    import socket
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        return s.connect_ex(("127.0.0.1", port)) == 0


def free_port(start: int = 5315) -> int:
    port = start
    while is_port_in_use(port):
        port += 1
    return port


def _topk(vec, k):
    from ..ops.topk import topk_dense
    return topk_dense(vec, k)
