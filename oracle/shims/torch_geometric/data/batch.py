from oracle.tp_nn import Batch  # noqa
