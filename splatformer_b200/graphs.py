"""Whole render steps replayed from a CUDA graph.

The fused path enqueues five kernels and two memsets per scene step plus a handful of small torch kernels; when a
step is short (few views per rank, small scenes) or the host is contended (eight ranks on one box), the Python and
launch cost per pass (about 0.55 ms, scripts/host_overhead.py) is what bounds the loop, not the device.  Capturing
the step -- forward and backward of every scene it renders -- into ONE CUDA graph makes a step cost the host a
single launch.  Collectives stay outside: an NCCL all-reduce issued while the stream captures invalidates the capture
(async work handle) or hangs (blocking form) on this stack (torch 2.11, NCCL 2.28.9), so a view-sharded step captures
its renders and issues `dist.sync_gaussian_grads`' reduces eagerly between replays, or is not captured at all.

What a captured region must not contain -- an event the host waits for, a flag read, a replay decision -- is exactly
what `RenderPlan.graph_safe` switches off.  So the recipe is: run the step eagerly a few times (the plan learns the
shapes' capacities and reads its flags as usual), then `capture_step`; after replays `plan.check()` tells whether any
captured pass overflowed, in which case that step has to be redone eagerly (the plan has learnt from it).
The tensors the step reads and writes are the ones that existed at capture time: new inputs are copied INTO them
(`param.data.copy_(...)`), gradients are read from the `.grad` tensors the captured backward created.
"""
from __future__ import annotations

from typing import Callable, Optional, Tuple

import torch

from .render import RenderPlan


def capture_step(step_fn: Callable[[int], None], plan: RenderPlan, passes: int
                 ) -> Tuple[Optional[torch.cuda.CUDAGraph], Optional[str]]:
    """Captures `step_fn(0)` -- which must render through `plan` -- into a CUDA graph.  `passes`: how many
    render_views calls the step makes (pinned landing slots are reserved for them up front: pinned memory cannot
    be allocated while a stream is capturing).  Returns (graph, None) or (None, reason)."""
    try:
        plan.graph_safe = True
        plan.graph_totals.clear()
        plan.reserve_graph_slots(2 * passes + 2)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            step_fn(0)                                # once more on a side stream, as torch's capture recipe asks
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        plan.graph_totals.clear()
        graph = torch.cuda.CUDAGraph()
        # thread_local: other threads of the process (a data loader, NCCL's watchdog) may keep calling the runtime
        with torch.cuda.graph(graph, capture_error_mode="thread_local"):
            step_fn(0)
        return graph, None
    except Exception as e:                            # noqa: BLE001  (the caller falls back to the eager step)
        plan.graph_safe = False
        torch.cuda.synchronize()
        return None, f"{type(e).__name__}: {str(e)[:200]}"
