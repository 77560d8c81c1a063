"""History-aggregation weights W0 (reference: cf_cpu/cf/behavior_aggregators.py:8-17):
nn.Linear(D, D, bias=False) initialised normal_(std=1e-2); its [D, D] weight is shared with cf_c."""
import torch
import torch.nn as nn

try:
    from .cpp_base import CPPBase
    from . import cf_c
except ImportError:
    from cpp_base import CPPBase
    import cf_c


class AggregatorWeights(CPPBase, nn.Module):
    def __init__(self, config):
        super().__init__()
        nn.Module.__init__(self)
        self.c_class = cf_c.modules.behavior_aggregators.AggregatorWeights
        self.f_c0 = nn.Linear(config.emb_dim, config.emb_dim, bias=False, dtype=torch.float32)
        nn.init.normal_(self.f_c0.weight, std=1e-2)
        self.aggregator_weights0 = self.f_c0.weight.detach().cpu().numpy()
        self.init_c_instance(aggregator_weights0=self.aggregator_weights0)
