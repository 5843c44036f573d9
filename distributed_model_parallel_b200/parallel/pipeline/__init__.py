from .p2p import (RecvForwardSendBackward, SendForwardRecvBackward, TensorMeta, backward_placeholder, recv_forward, recv_tensor,  # noqa: F401
                  send_forward, send_tensor)
from .partition import layer_costs, measured_boundaries, partition_balanced, reference_split, split_sequential  # noqa: F401
from .schedule import PipelineStage, generate_schedule, max_in_flight, simulate_schedule  # noqa: F401
