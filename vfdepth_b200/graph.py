"""CUDA-graph capture of a training step built from the drop-in modules.

The step of this path is ~20 kernels of 0.01-3 ms plus ~60 tiny torch launches (pose products, loss combination,
autograd glue) whose launch gaps cost about 0.4 ms of an 11 ms step.  All ABI entry points are capture-safe: they only
enqueue kernels / memsets on the stream they are given, never synchronise, never allocate, and take every size from
device memory (the compact voxel lists' lengths stay on the device).  ``capture`` records one call of ``fn`` (forward and
backward, inputs read from fixed tensors) into a ``torch.cuda.CUDAGraph`` and returns a replay function.
"""
import torch


def capture(fn, warmup=3):
    """-> (replay, outputs).  ``fn()`` must read its inputs from tensors that stay alive and in place; ``outputs`` is
    what the captured call returned (refreshed in place by every ``replay()``)."""
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):            # warm-up on a side stream, as graph capture requires
        for _ in range(warmup):
            fn()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        outputs = fn()
    return graph.replay, outputs
