"""Device-resident replacement for the reference's input pipeline.

ref: ray-jobs/pytorch_llm_ray.py:191-216 builds `TextDataset(token_ids, seq_len)` (stride-1 windows, targets shifted by
one, :107-119), optionally `Subset(first 16000)` (:198-201), a shuffling `DataLoader(batch_size, drop_last=True)` and
hands it to `train.torch.prepare_data_loader`, which adds a `DistributedSampler` and moves every batch to the device.
Here the token stream lives in HBM once, the SAME sampler decides which windows each rank sees (so the two pipelines
are comparable batch for batch), and each batch is produced by one gather kernel instead of B Python `__getitem__`
calls, a collate and a pinned-memory copy.
"""
import torch
from torch.utils.data import DistributedSampler

from ._lib import B2Error, call, ptr, stream_ptr


class _Windows:
    """Index space of TextDataset: window i covers tokens[i : i + seq_len + 1]."""

    def __init__(self, n_tokens, seq_len, limit=None):
        self.n = max(0, n_tokens - seq_len - 1)            # ref :111 — keeps the shifted target in bounds
        if limit is not None:
            self.n = min(self.n, limit)                    # ref :198-201 — test_run keeps the first 16000 windows

    def __len__(self):
        return self.n


class ResidentWindowLoader:
    """Iterates (inputs, targets) int64 [batch_size, seq_len] device tensors, like the prepared DataLoader of the
    reference. `sampler` is a stock torch DistributedSampler (exposed so `loader.sampler.set_epoch(e)` works, ref :265)."""

    def __init__(self, token_ids, seq_len, batch_size, device, shuffle=True, drop_last=True, num_replicas=1, rank=0,
                 seed=0, limit=None):
        if token_ids.dtype != torch.int64 or token_ids.dim() != 1:
            raise B2Error("token_ids must be a 1-D int64 tensor (the reference's train.ids.pt)")
        self.tokens = token_ids.to(device).contiguous()
        self.seq_len, self.batch_size, self.device, self.drop_last = int(seq_len), int(batch_size), device, drop_last
        self.dataset = _Windows(self.tokens.numel(), seq_len, limit)
        if len(self.dataset) == 0:
            raise ValueError("Dataset empty after loading")          # same failure as ref :203-204
        self.sampler = DistributedSampler(self.dataset, num_replicas=num_replicas, rank=rank, shuffle=shuffle, seed=seed,
                                          drop_last=False)
        self._err = torch.zeros(1, dtype=torch.int32, device=device)

    def __len__(self):
        n = len(self.sampler)
        return n // self.batch_size if self.drop_last else (n + self.batch_size - 1) // self.batch_size

    def gather(self, starts):
        """One batch for the given window start indices (int64 device tensor [B])."""
        B = starts.numel()
        x = torch.empty(B, self.seq_len, dtype=torch.int64, device=self.device)
        y = torch.empty(B, self.seq_len, dtype=torch.int64, device=self.device)
        with torch.cuda.device(self.device):
            call("b2_gather_windows", ptr(self.tokens), self.tokens.numel(), ptr(starts), B, self.seq_len, ptr(x), ptr(y),
                 ptr(self._err), stream_ptr())
        return x, y

    def __iter__(self):
        idx = list(iter(self.sampler))
        # the kernel flags a window that does not fit the token stream and zero-fills it; the reference's __getitem__
        # would hand collate a short tensor and fail there. Window starts come from the sampler, so they are checked on
        # the host, before anything is gathered — no device flag to poll.
        if idx and (min(idx) < 0 or max(idx) + self.seq_len + 1 > self.tokens.numel()):
            raise IndexError(f"window start outside [0, {self.tokens.numel() - self.seq_len - 1}] for seq_len {self.seq_len}")
        order = torch.tensor(idx, dtype=torch.int64).to(self.device, non_blocking=True)
        for i in range(len(self)):
            yield self.gather(order[i * self.batch_size:(i + 1) * self.batch_size])

    def check(self):
        """True if no gather since construction met an out-of-range window (one host sync; for callers of gather())."""
        return int(self._err.item()) == 0
