"""Python mirror of the training configuration (reference: cf_cpu/cf/cf_config.py:6-40).

Same constructor keywords and attributes; `init_c_instance` builds cf_c.modules.CFConfig with the
reference's 13 keywords (including the misspelt `num_subepoches`, init_modules.cpp:26).  The GPU
knobs (`seed`, `update_mode`, `max_blocks`, ...) are optional attributes copied onto the C++ object
when present, so every existing YAML stays valid.
"""
try:
    from .cpp_base import CPPBase
    from . import cf_c
except ImportError:  # run as a script from this directory, like the reference
    from cpp_base import CPPBase
    import cf_c

GPU_KNOBS = ("seed", "sample_offset", "chunk", "deterministic", "update_mode", "warps_per_block", "max_blocks",
             "sync_average", "sync_grads", "w0_tail_mode", "eval_precision", "history_mode", "engine", "tile_sampler", "exchange_mode")


class CFConfig(CPPBase):
    def __init__(self, emb_dim=64, num_negs=4, max_his=8, num_users=128, num_items=128, train_size=128,
                 neg_sampler=0, tile_size=1024, num_subepoches=2, refresh_interval=2048, l2=1.e-3, clip_val=0.1,
                 milestones=None, l_r=1.e-3, **gpu_knobs):
        super().__init__()
        self.c_class = cf_c.modules.CFConfig
        self.emb_dim, self.num_negs = emb_dim, num_negs
        self.num_users, self.num_items, self.train_size = num_users, num_items, train_size
        self.neg_sampler, self.tile_size, self.refresh_interval = neg_sampler, tile_size, refresh_interval
        self.num_subepoches = num_subepoches
        self.l2, self.clip_val, self.l_r = l2, clip_val, l_r
        self.milestones = list(milestones) if milestones else [0]
        self.en_his = True
        self.max_his = max_his
        self.seed = 2022
        for k, v in gpu_knobs.items():
            if k not in GPU_KNOBS:
                raise TypeError(f"unknown CFConfig argument {k!r}")
            setattr(self, k, v)

    def init_c_instance(self):
        self.c_instance = self.c_class(
            emb_dim=int(self.emb_dim), num_negs=int(self.num_negs), num_users=int(self.num_users),
            num_items=int(self.num_items), train_size=int(self.train_size), neg_sampler=int(self.neg_sampler),
            tile_size=int(self.tile_size), refresh_interval=int(self.refresh_interval),
            num_subepoches=int(self.num_subepoches), l2=float(self.l2), clip_val=float(self.clip_val),
            milestones=[int(m) for m in self.milestones], l_r=float(self.l_r))
        for k in GPU_KNOBS:
            if hasattr(self, k):
                setattr(self.c_instance, k, int(getattr(self, k)))
