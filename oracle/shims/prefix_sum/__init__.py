def prefix_sum_cuda(*a, **k):
    raise NotImplementedError
