"""Checkpoint helpers: torch modules / tensors as rabit global and local models.

``rabit.checkpoint`` pickles arbitrary objects (reference python/rabit.py:279-351); for torch state
that would pickle CUDA storages, so these helpers move tensors to the CPU first and restore them onto
the module's devices.
"""
import torch

from .. import rabit


def _to_cpu(obj):
    if torch.is_tensor(obj):
        return obj.detach().cpu()
    if isinstance(obj, dict):
        return {k: _to_cpu(v) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return type(obj)(_to_cpu(v) for v in obj)
    return obj


def save(module_or_state, local_state=None, extra=None):
    """Commit a new version: global = module.state_dict() (identical on all ranks), local = per-rank state."""
    state = module_or_state.state_dict() if hasattr(module_or_state, "state_dict") else module_or_state
    rabit.checkpoint({"state": _to_cpu(state), "extra": extra}, None if local_state is None else _to_cpu(local_state))


def load(module=None, with_local=False):
    """Returns (version, extra[, local]); loads the state into `module` when a checkpoint exists."""
    if with_local:
        version, blob, local = rabit.load_checkpoint(with_local=True)
    else:
        version, blob = rabit.load_checkpoint()
        local = None
    extra = None
    if version != 0 and blob is not None:
        extra = blob.get("extra")
        if module is not None:
            module.load_state_dict(blob["state"])
    return (version, extra, local) if with_local else (version, extra)
