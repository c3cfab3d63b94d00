from oracle.tp_ops import global_add_pool  # noqa
