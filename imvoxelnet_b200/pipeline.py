"""Host-buffer front end of the hot path: copies and compute on separate streams.

``StreamedHotPath`` is what a serving loop calls when its feature maps live in
(pinned) host memory: every ``run`` step uploads the next batch, computes the
current one and downloads the previous result concurrently (one copy-in
stream, the caller's compute stream, one copy-out stream; double-buffered
device inputs and host outputs).  The reference has nothing like it -- it
moves tensors with ``.to(device)`` on the default stream
(mmdet3d/models/detectors/imvoxelnet.py:61,66) and returns device tensors.
"""
import torch

__all__ = ["StreamedHotPath", "GraphedHotPath"]


class StreamedHotPath:
    def __init__(self, compute, depth=2):
        """``compute(features_dev, projection_dev) -> (list of tensors, valid)`` is
        the device-side step (e.g. back-projection + neck)."""
        self.compute = compute
        self.depth = depth
        self.copy_in = torch.cuda.Stream()
        self.copy_out = torch.cuda.Stream()
        self._dev_in = [None] * depth        # (features, projection) device buffers
        self._host_out = [None] * depth      # pinned host outputs
        self._in_ready = [torch.cuda.Event() for _ in range(depth)]
        self._computed = [torch.cuda.Event() for _ in range(depth)]
        self._out_done = [torch.cuda.Event() for _ in range(depth)]
        self._used = [False] * depth
        self._pending = []                   # device outputs kept alive until downloaded

    @staticmethod
    def _pinned_like(t):
        flat = torch.empty(t.numel(), dtype=t.dtype).pin_memory()
        return flat.as_strided(t.shape, t.stride())

    def run(self, batches):
        """``batches``: iterable of (features, projection) pinned host tensors.
        Yields, in order, ``(list of pinned host outputs, pinned valid)`` for
        every batch; a yielded result is valid until ``depth`` more batches
        have been consumed from the generator."""
        compute_stream = torch.cuda.current_stream()
        inflight = []
        for i, (feats, proj) in enumerate(batches):
            slot = i % self.depth
            # upload: the slot's device buffers are free once its previous compute is done
            with torch.cuda.stream(self.copy_in):
                if self._used[slot]:
                    self.copy_in.wait_event(self._computed[slot])
                if self._dev_in[slot] is None:
                    self._dev_in[slot] = (torch.empty_strided(feats.shape, feats.stride(), dtype=feats.dtype,
                                                              device="cuda"),
                                          torch.empty_like(proj, device="cuda"))
                self._dev_in[slot][0].copy_(feats, non_blocking=True)
                self._dev_in[slot][1].copy_(proj, non_blocking=True)
                self._in_ready[slot].record(self.copy_in)
            # compute
            compute_stream.wait_event(self._in_ready[slot])
            outs, valid = self.compute(*self._dev_in[slot])
            self._computed[slot].record(compute_stream)
            # download: the slot's host buffers are free once the consumer has moved on
            with torch.cuda.stream(self.copy_out):
                self.copy_out.wait_event(self._computed[slot])
                if self._host_out[slot] is None:
                    self._host_out[slot] = ([self._pinned_like(o) for o in outs], self._pinned_like(valid))
                for h, o in zip(self._host_out[slot][0], outs):
                    h.copy_(o, non_blocking=True)
                    o.record_stream(self.copy_out)
                self._host_out[slot][1].copy_(valid, non_blocking=True)
                valid.record_stream(self.copy_out)
                self._out_done[slot].record(self.copy_out)
            self._used[slot] = True
            inflight.append(slot)
            if len(inflight) == self.depth:          # oldest result must be handed out before its slot is reused
                done = inflight.pop(0)
                self._out_done[done].synchronize()
                yield self._host_out[done]
        for done in inflight:
            self._out_done[done].synchronize()
            yield self._host_out[done]


class GraphedHotPath:
    """Replays a device-side step as a CUDA graph (one graph per input signature).

    A forward of a small scene is a dozen kernels of tens of microseconds each:
    issued eagerly from Python the launch gaps cost more than the kernels.  The
    first call with a given (shape, dtype, strides) signature warms up, captures
    ``compute(*inputs)`` into a graph with static input buffers, and later calls
    copy their inputs into those buffers and replay.  Outputs are the graph's
    static output tensors: consume (or clone) them before the next call.
    Host-side values baked in at capture (``hw``, origins, projection device
    tensors passed as inputs are fine) must not change between calls.
    """

    def __init__(self, compute, warmup=2):
        self.compute = compute
        self.warmup = warmup
        self._graphs = {}

    @staticmethod
    def _signature(tensors):
        return tuple((tuple(t.shape), t.dtype, tuple(t.stride()), t.device.index) for t in tensors)

    def __call__(self, *inputs):
        key = self._signature(inputs)
        entry = self._graphs.get(key)
        if entry is None:
            static_in = [torch.empty_strided(t.shape, t.stride(), dtype=t.dtype, device=t.device) for t in inputs]
            for s, t in zip(static_in, inputs):
                s.copy_(t)
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side), torch.no_grad():
                for _ in range(self.warmup):
                    self.compute(*static_in)
            torch.cuda.current_stream().wait_stream(side)
            graph = torch.cuda.CUDAGraph()
            with torch.no_grad(), torch.cuda.graph(graph):
                static_out = self.compute(*static_in)
            entry = self._graphs[key] = (graph, static_in, static_out)
        graph, static_in, static_out = entry
        for s, t in zip(static_in, inputs):
            if s.data_ptr() != t.data_ptr():
                s.copy_(t)
        graph.replay()
        return static_out
