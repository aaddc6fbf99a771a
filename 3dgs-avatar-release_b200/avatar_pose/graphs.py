"""CUDA-graph capture of a whole hot-path step (posing -> rasterize -> loss -> backward).

At 512x512 / 50k Gaussians the GPU work of one training step is ~0.4 ms while the Python / autograd / launch overhead of
the ~35 launches is larger, so the step is host-bound when run eagerly.  `GraphedStep` captures the step once and replays
it with a single launch.  Inputs that change per frame (pose, camera, target image) live in static device tensors that
the caller overwrites before `replay()`; outputs are static tensors too.

The rasterizer's pair capacity is fixed inside a graph: `replay(check=True)` synchronises, reads the pair count the
graph copied into pinned memory and re-captures with a larger capacity if it overflowed (the frame is then re-run).
"""
import torch

from diff_gaussian_rasterization import _C as _gs


class GraphedStep:
    def __init__(self, fn, warmup=3):
        """fn(): runs one step reading its per-frame inputs from tensors whose storage stays fixed; returns a tensor or
        a tuple of tensors (kept as static outputs)."""
        self.fn, self.warmup = fn, warmup
        self.graph, self.out = None, None
        self.capture()

    def capture(self):
        cur = torch.cuda.current_stream()
        side = torch.cuda.Stream()
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            for _ in range(self.warmup):  # eager warm-up: sizes the pair capacity and the scratch buffers
                self.fn()
        cur.wait_stream(side)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.out = self.fn()
        return self

    def replay(self, check=False):
        self.graph.replay()
        if check:
            torch.cuda.current_stream().synchronize()
            n, over = _gs.graph_overflow()
            if over:  # captured capacity too small for this frame: re-capture (capacity already raised) and re-run
                self.capture()
                self.graph.replay()
                torch.cuda.current_stream().synchronize()
        return self.out

    def release(self):
        """Frees the captured graph and its static outputs (call before tearing a process group / the CUDA context down:
        a live graph keeps kernels of the communicator and pointers into pool memory)."""
        self.out = None
        if self.graph is not None:
            self.graph.reset()
            self.graph = None
