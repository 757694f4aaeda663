"""Keeping one GPU busy across the host synchronisations of the matcher forward.

``OnePosePlus_model.forward`` has to learn the number of coarse matches on the host before it can size the
fine-level tensors (the reference does the same through ``torch.where``: utils/coarse_matching.py:196-204), so
one forward leaves the GPU idle twice: while the host wakes up after that read-back, and between forwards.  The
reference hides this by running several Ray workers per GPU (``n_gpus_per_worker`` < 1 in
src/inference/inference_OnePosePlus.py:67-100, src/KeypointFreeSfM/coarse_match/coarse_match.py:51-93).  Here:
one process per GPU, ``n_streams`` host threads, each with its own CUDA stream (and therefore its own
workspaces, see ``ops.workspaces``), pulling independent units (frame batches / image pairs) from one queue.
While one thread waits for its match count, the other thread's kernels run.

    pipe = StreamPipeline(2)
    results = pipe.map(lambda frames: run_forward(model, frames), batches)   # results in input order

``map`` orders the worker streams after the caller's current stream on entry and the caller's stream after the
workers on exit, so it can be bracketed by CUDA events on the caller's stream like any other stream work.
Tensors returned by ``fn`` were allocated on a worker stream: call ``t.record_stream(torch.cuda.current_stream())``
before consuming them on the caller's stream if they may be freed while that work is still queued.
"""
import queue
import threading
from typing import Callable, Iterable, List

import torch


class StreamPipeline:
    def __init__(self, n_streams: int = 2, device=None):
        if n_streams < 1:
            raise ValueError("n_streams must be >= 1")
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.n_streams = n_streams
        self.streams = [torch.cuda.Stream(device=self.device) for _ in range(n_streams)]

    def warm(self, fn: Callable) -> None:
        """Run ``fn()`` once on every worker stream, one after the other, from the calling thread (builds per-stream
        state such as workspaces and CUDA graphs outside any timed region)."""
        for s in self.streams if self.n_streams > 1 else []:
            s.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(s), torch.no_grad():
                fn()
            s.synchronize()
        if self.n_streams == 1:
            with torch.no_grad():
                fn()
            torch.cuda.current_stream(self.device).synchronize()

    def map(self, fn: Callable, items: Iterable) -> List:
        items = list(items)
        results = [None] * len(items)
        if not items:
            return results
        caller = torch.cuda.current_stream(self.device)
        if self.n_streams == 1:  # no threads: run on the caller's stream
            for i, it in enumerate(items):
                results[i] = fn(it)
            return results
        start = torch.cuda.Event()
        start.record(caller)
        todo = queue.SimpleQueue()
        for i, it in enumerate(items):
            todo.put((i, it))
        errors = []
        done_events = [None] * self.n_streams

        def worker(w):
            try:
                torch.cuda.set_device(self.device)
                s = self.streams[w]
                s.wait_event(start)
                with torch.cuda.stream(s), torch.no_grad():
                    while not errors:
                        try:
                            i, it = todo.get_nowait()
                        except queue.Empty:
                            break
                        results[i] = fn(it)
                    ev = torch.cuda.Event()
                    ev.record(s)
                    done_events[w] = ev
            except BaseException as e:  # surfaced in the caller
                errors.append(e)

        threads = [threading.Thread(target=worker, args=(w,), daemon=True) for w in range(self.n_streams)]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        if errors:
            raise errors[0]
        for ev in done_events:
            if ev is not None:
                caller.wait_event(ev)
        return results
