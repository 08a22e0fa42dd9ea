"""Whole-pass CUDA graphs for launch-bound graphs (Cora / PubMed sized: ~130 kernel launches and a few
milliseconds of Python per training pass, against ~3 ms of kernel time).

`GraphedPass(fn, inputs)` warms `fn` up on a side stream, captures ONE call - forward integration, adjoint
backward, every kernel of libgrandb200 and the torch glue between them - into a `torch.cuda.CUDAGraph`, and
replays it with new inputs copied into the captured buffers.  Nothing in the hot path synchronises the device
(plans, weight concatenations and function attributes are cached on the host before capture), which is what
makes the capture legal.  The graph topology is fixed: use it with the fixed-grid solvers (euler / midpoint /
rk4), an unchanged graph and unchanged shapes; adaptive solvers decide their step count on the host.
"""
import torch


class GraphedPass:
    def __init__(self, fn, inputs, warmup=3):
        if not all(t.is_cuda for t in inputs):
            raise RuntimeError('GraphedPass needs CUDA tensors')
        self.fn = fn
        self.static_in = [t.detach().clone() for t in inputs]
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(warmup):
                fn(*self.static_in)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.static_out = fn(*self.static_in)

    def __call__(self, *inputs):
        for dst, src in zip(self.static_in, inputs):
            if dst.data_ptr() != src.data_ptr():
                dst.copy_(src, non_blocking=True)
        self.graph.replay()
        return self.static_out
