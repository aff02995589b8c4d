"""Slide-at-a-time inference with the host->device copy of the next bag hidden behind the forward of the current one.

The reference's eval loop hands one slide per call to the model (`ModelInterface.step`, code/models/model_interface.py:
316-324: `logits = self(x.contiguous())`), with the features arriving from a DataLoader in host memory
(code/datasets/feature_dataloader.py:252-254).  On a B200 the forward of a 50k x 1024 bag takes ~1 ms while its 205 MB
cross PCIe in ~3.7 ms, so the loop is copy-bound; `StreamedInference` keeps `depth` device staging buffers and a copy
stream so that, in steady state, a slide costs max(copy, compute) instead of their sum.  Arithmetic is untouched: every
bag goes through `TransMIL.forward` (the C ABI) exactly as a direct call would.
"""
import torch

__all__ = ["StreamedInference"]


class StreamedInference:
    def __init__(self, model, depth=2):
        if model.training:
            raise ValueError("StreamedInference is an eval-mode loop (model.eval())")
        self.model = model
        self.device = model.cls_token.device
        if self.device.type != "cuda":
            raise RuntimeError("StreamedInference needs the model on a CUDA device; there is no CPU path")
        self.depth = max(2, int(depth))
        self.copy_stream = torch.cuda.Stream(self.device)
        self._slots = [None] * self.depth

    def _stage(self, i, x):
        """Start the H2D copy of bag `x` (host, ideally pinned) into slot i on the copy stream."""
        if x.dim() == 2:
            x = x.unsqueeze(0)
        slot = self._slots[i]
        nbytes = x.numel() * x.element_size()              # slots are raw bytes: fp32 and fp16 feature bags share them
        if slot is None or slot["buf"].numel() < nbytes:
            slot = self._slots[i] = dict(buf=torch.empty(nbytes, dtype=torch.uint8, device=self.device),
                                         ready=torch.cuda.Event(), free=torch.cuda.Event())
            slot["buf"].record_stream(self.copy_stream)
            slot["free"].record(torch.cuda.current_stream(self.device))
        self.copy_stream.wait_event(slot["free"])          # the forward that last read this slot is done
        dst = slot["buf"][:nbytes].view(x.dtype).view(x.shape)
        with torch.cuda.stream(self.copy_stream):
            dst.copy_(x, non_blocking=True)
            slot["ready"].record(self.copy_stream)
        return dst

    @torch.no_grad()
    def __call__(self, bags):
        """bags: iterable of host tensors [n_i, D] / [1, n_i, D] (fp32, or fp16 feature files as stored -- the model reads halves
        directly; pinned memory makes the copy asynchronous).
        Yields the logits of each bag as a host tensor [1, n_classes], in order."""
        main = torch.cuda.current_stream(self.device)
        it = iter(bags)
        pending = []                                        # (slot index, device view)
        nxt = 0
        for _ in range(self.depth - 1):                     # prime the pipeline
            try:
                x = next(it)
            except StopIteration:
                break
            pending.append((nxt, self._stage(nxt, x)))
            nxt = (nxt + 1) % self.depth
        prev = None                                         # (logits_host, event) of the previous bag
        while pending:
            try:                                            # issue the copy of the bag after the one about to run
                x = next(it)
                pending.append((nxt, self._stage(nxt, x)))
                nxt = (nxt + 1) % self.depth
            except StopIteration:
                pass
            i, xd = pending.pop(0)
            slot = self._slots[i]
            main.wait_event(slot["ready"])
            logits = self.model(xd)
            slot["free"].record(main)
            host = torch.empty(logits.shape, dtype=logits.dtype, pin_memory=True)
            host.copy_(logits, non_blocking=True)
            done = torch.cuda.Event()
            done.record(main)
            if prev is not None:
                prev[1].synchronize()
                yield prev[0]
            prev = (host, done)
        if prev is not None:
            prev[1].synchronize()
            yield prev[0]
