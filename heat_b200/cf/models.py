"""Embedding tables (reference: cf_cpu/cf/models.py:8-32): torch nn.Embedding user/item tables,
normal_(std=1e-2) init, handed to cf_c as numpy views that SHARE storage with the torch
parameters -- after Engine.train_one_epoch() they hold the trained values."""
import torch
import torch.nn as nn

try:
    from .cpp_base import CPPBase
    from . import cf_c
except ImportError:
    from cpp_base import CPPBase
    import cf_c


class Model(CPPBase, nn.Module):
    def __init__(self, config):
        super().__init__()
        nn.Module.__init__(self)
        self.user_embedding = nn.Embedding(config.num_users, config.emb_dim, dtype=torch.float32)
        self.item_embedding = nn.Embedding(config.num_items, config.emb_dim, dtype=torch.float32)
        nn.init.normal_(self.user_embedding.weight, std=1e-2)
        nn.init.normal_(self.item_embedding.weight, std=1e-2)
        self.user_weights = None
        self.item_weights = None


class MatrixFactorization(Model):
    def __init__(self, config):
        super().__init__(config)
        self.c_class = cf_c.modules.models.MatrixFactorization

    def init_c_instance(self, config=None):
        self.user_weights = self.user_embedding.weight.detach().cpu().numpy()
        self.item_weights = self.item_embedding.weight.detach().cpu().numpy()
        self.c_instance = self.c_class(cf_config=config.c_instance, user_weights=self.user_weights,
                                       item_weights=self.item_weights)
