from . import functional
from .functional import conv_bn_act, cross_entropy, global_avgpool_flatten, linear, path_counts
