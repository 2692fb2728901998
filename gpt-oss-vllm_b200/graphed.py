"""CUDA-graph wrapper for a decode step: the way a serving engine runs this path.

vLLM replays its decode steps from CUDA graphs unless `--enforce-eager` is given (V/.../moe_runner.py); the C ABI is
written for that (no allocation, no host sync, no data-dependent host control flow).  `GraphedForward` is the same thing as a
small public class: it captures `fn(*device_inputs) -> device_output` once over static buffers and then serves calls that bring
HOST (pinned) inputs: H2D copy into the static buffers, one graph replay, D2H copy of the result -- all on one stream.
"""
from __future__ import annotations

from typing import Callable, Sequence

import torch


class GraphedForward:
    def __init__(self, fn: Callable[..., torch.Tensor], example_inputs: Sequence[torch.Tensor], warmup: int = 2):
        if not example_inputs or any(t.device.type != "cuda" for t in example_inputs):
            raise RuntimeError("GraphedForward needs CUDA example inputs: there is no CPU fallback on this path")
        self.device = example_inputs[0].device
        self.inputs = [torch.empty_like(t) for t in example_inputs]          # static device buffers the graph reads
        for dst, src in zip(self.inputs, example_inputs):
            dst.copy_(src)
        self.stream = torch.cuda.Stream(device=self.device)
        self.graph = torch.cuda.CUDAGraph()
        torch.cuda.synchronize(self.device)
        with torch.cuda.stream(self.stream):
            for _ in range(max(1, warmup)):                                   # lazy initialisation outside the capture
                fn(*self.inputs)
            self.stream.synchronize()
            with torch.cuda.graph(self.graph, stream=self.stream):
                self.output = fn(*self.inputs)
        torch.cuda.synchronize(self.device)

    def replay(self) -> torch.Tensor:
        """Device-resident inputs already in `self.inputs`: replay on the current stream."""
        self.graph.replay()
        return self.output

    def __call__(self, *host_inputs: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
        """host (pinned) inputs -> host output; synchronises the stream before returning."""
        if len(host_inputs) != len(self.inputs):
            raise ValueError(f"expected {len(self.inputs)} inputs, got {len(host_inputs)}")
        with torch.cuda.stream(self.stream):
            for dst, src in zip(self.inputs, host_inputs):
                if src.shape != dst.shape or src.dtype != dst.dtype:
                    raise ValueError(f"input {tuple(src.shape)}/{src.dtype} does not match the captured {tuple(dst.shape)}/{dst.dtype}")
                dst.copy_(src, non_blocking=True)
            self.graph.replay()
            if out is None:
                out = torch.empty(self.output.shape, dtype=self.output.dtype, pin_memory=True)
            out.copy_(self.output, non_blocking=True)
            self.stream.synchronize()
        return out
