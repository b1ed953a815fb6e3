"""Checkpoints in the reference's format, written without stalling the training stream.

Mirror of /root/reference/audio_slowfast/utils/checkpoint.py: directory / file naming (:17-52), ``get_last_checkpoint``
/ ``has_checkpoint`` (:55-80), ``is_checkpoint_epoch`` (:83-93), ``save_checkpoint`` (:96-125: ``torch.save({"epoch",
"model_state", "optimizer_state", "cfg"})``, master process only, DDP wrapper stripped) and ``load_checkpoint``
(:128-200: CPU load, optional renaming / exclusion, keep the entries whose name AND shape match, optimizer state unless
``epoch_reset``).  The ``state_dict`` layout is the reference's (664 fp32 OIHW entries for AudioSlowFast), so files move
between the two implementations unchanged.

``AsyncCheckpointWriter`` is the B200-side addition: ``save_checkpoint`` in the reference blocks the training loop for
the device->host copy of 107 MB of weights + 107 MB of momentum and for the file write.  Here the state is copied into
pinned host buffers on a side stream (the training stream only waits for an event BEFORE its next optimizer step
overwrites the weights) and ``torch.save`` runs in a background thread.
"""
from __future__ import annotations

import os
import threading
from collections import OrderedDict

import torch


def make_checkpoint_dir(path_to_job):
    d = os.path.join(path_to_job, "checkpoints")
    os.makedirs(d, exist_ok=True)
    return d


def get_checkpoint_dir(path_to_job):
    return os.path.join(path_to_job, "checkpoints")


def get_path_to_checkpoint(path_to_job, epoch, is_best_epoch=False):
    name = "checkpoint_best.pyth" if is_best_epoch else "checkpoint_epoch_{:05d}.pyth".format(epoch)
    return os.path.join(get_checkpoint_dir(path_to_job), name)


def get_last_checkpoint(path_to_job):
    d = get_checkpoint_dir(path_to_job)
    names = [f for f in (os.listdir(d) if os.path.exists(d) else []) if "checkpoint" in f]
    assert len(names), "No checkpoints found in '{}'.".format(d)
    return os.path.join(d, sorted(names)[-1])


def has_checkpoint(path_to_job):
    d = get_checkpoint_dir(path_to_job)
    return any("checkpoint" in f for f in (os.listdir(d) if os.path.exists(d) else []))


def is_checkpoint_epoch(cfg, cur_epoch):
    if cur_epoch + 1 == cfg.SOLVER.MAX_EPOCH:
        return True
    return (cur_epoch + 1) % cfg.TRAIN.CHECKPOINT_PERIOD == 0


def _is_master():
    import torch.distributed as dist

    return not (dist.is_available() and dist.is_initialized()) or dist.get_rank() == 0


def _unwrap(model):
    return model.module if hasattr(model, "module") else model


def _to_host(obj, pinned, key=""):
    """Deep copy of a (nested) state with every CUDA tensor copied asynchronously into a pinned host tensor."""
    if torch.is_tensor(obj):
        if not obj.is_cuda:
            return obj.detach().clone()
        buf = pinned.get(key)
        if buf is None or buf.shape != obj.shape or buf.dtype != obj.dtype:
            buf = pinned[key] = torch.empty(obj.shape, dtype=obj.dtype, pin_memory=True)
        buf.copy_(obj.detach(), non_blocking=True)
        return buf
    if isinstance(obj, dict):
        return type(obj)((k, _to_host(v, pinned, "%s/%s" % (key, k))) for k, v in obj.items())
    if isinstance(obj, (list, tuple)):
        return type(obj)(_to_host(v, pinned, "%s/%d" % (key, i)) for i, v in enumerate(obj))
    return obj


class AsyncCheckpointWriter:
    """``writer.save(path_to_job, model, optimizer, epoch, cfg)`` returns as soon as the device->host copies are queued on
    the writer's stream; the file is written by a background thread.  ``fence()`` (call it before the next optimizer
    step) makes the current stream wait until the snapshot has left the device; ``wait()`` joins the file write."""

    def __init__(self, device=None):
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.stream = torch.cuda.Stream(device=self.device)
        self._pinned = {}
        self._thread = None
        self._copied = None
        self.last_path = None
        self.error = None

    def save(self, path_to_job, model, optimizer, epoch, cfg, is_best_epoch=False):
        if not _is_master():
            return None
        self.wait()  # one snapshot in flight: the pinned buffers are reused
        make_checkpoint_dir(path_to_job)
        path = get_path_to_checkpoint(path_to_job, epoch + 1, is_best_epoch)
        self.stream.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(self.stream):
            ckpt = {"epoch": epoch,
                    "model_state": _to_host(_unwrap(model).state_dict(), self._pinned, "m"),
                    "optimizer_state": _to_host(optimizer.state_dict(), self._pinned, "o") if optimizer is not None else {},
                    "cfg": cfg.dump() if hasattr(cfg, "dump") else cfg}
            self._copied = torch.cuda.Event()
            self._copied.record(self.stream)
        copied = self._copied

        def _write():
            try:
                copied.synchronize()
                # the pinned buffers are reused by the next snapshot: serialise private copies
                plain = {"epoch": ckpt["epoch"], "cfg": ckpt["cfg"],
                         "model_state": OrderedDict((k, v.clone() if torch.is_tensor(v) else v)
                                                    for k, v in ckpt["model_state"].items()),
                         "optimizer_state": _clone_tree(ckpt["optimizer_state"])}
                tmp = path + ".tmp"
                with open(tmp, "wb") as f:
                    torch.save(plain, f)
                os.replace(tmp, path)
            except Exception as e:  # surfaced by wait()
                self.error = e

        self._thread = threading.Thread(target=_write, daemon=True)
        self._thread.start()
        self.last_path = path
        return path

    def fence(self):
        """The current stream waits until the last snapshot's device->host copies are done (call before the weights
        are modified again)."""
        if self._copied is not None:
            torch.cuda.current_stream(self.device).wait_event(self._copied)

    def wait(self):
        if self._thread is not None:
            self._thread.join()
            self._thread = None
        if self.error is not None:
            e, self.error = self.error, None
            raise e


def _clone_tree(obj):
    if torch.is_tensor(obj):
        return obj.clone()
    if isinstance(obj, dict):
        return type(obj)((k, _clone_tree(v)) for k, v in obj.items())
    if isinstance(obj, (list, tuple)):
        return type(obj)(_clone_tree(v) for v in obj)
    return obj


def save_checkpoint(path_to_job, model, optimizer, epoch, cfg, is_best_epoch=False):
    """checkpoint.py:96-125, synchronous (same file as the reference writes)."""
    if not _is_master():
        return None
    make_checkpoint_dir(path_to_job)
    sd = _unwrap(model).state_dict()
    checkpoint = {"epoch": epoch, "model_state": OrderedDict((k, v.detach().cpu()) for k, v in sd.items()),
                  "optimizer_state": optimizer.state_dict() if optimizer is not None else {},
                  "cfg": cfg.dump() if hasattr(cfg, "dump") else cfg}
    path = get_path_to_checkpoint(path_to_job, epoch + 1, is_best_epoch)
    with open(path, "wb") as f:
        torch.save(checkpoint, f)
    return path


def load_checkpoint(path_to_checkpoint, model, data_parallel=True, optimizer=None, epoch_reset=False,
                    clear_name_pattern=(), exclude_layers=()):
    """checkpoint.py:128-200.  Returns the epoch stored in the file (-1 when absent or ``epoch_reset``)."""
    assert os.path.exists(path_to_checkpoint), "Checkpoint '{}' not found".format(path_to_checkpoint)
    ms = model.module if (data_parallel and hasattr(model, "module")) else model
    with open(path_to_checkpoint, "rb") as f:
        checkpoint = torch.load(f, map_location="cpu")
    state = checkpoint["model_state"]
    for item in clear_name_pattern:
        state = OrderedDict((k.replace(item, "") if item in k else k, v) for k, v in state.items())
    model_dict = ms.state_dict()
    state = {k: v for k, v in state.items() if not any(ex in k for ex in exclude_layers)}
    match = {k: v for k, v in state.items() if k in model_dict and v.size() == model_dict[k].size()}
    not_loaded = [k for k in model_dict if k not in match]
    ms.load_state_dict(match, strict=False)
    if hasattr(ms, "engine"):
        ms.engine.invalidate_weights()
    if "epoch" in checkpoint and not epoch_reset:
        epoch = checkpoint["epoch"]
        if optimizer is not None and checkpoint.get("optimizer_state"):
            optimizer.load_state_dict(checkpoint["optimizer_state"])
    else:
        epoch = -1
    load_checkpoint.not_loaded = not_loaded
    return epoch
