"""The reference's model plugin mechanism, re-expressed without Lightning.

ModelInterface.load_model / instancialize (code/models/model_interface.py:1256-1293):
`cfg.Model.name` -> `importlib.import_module(f'models.{name}')`, class name = name
(CamelCase if it contains '_'), constructor kwargs = intersection of the class'
__init__ argument names with the keys of cfg.Model.  `inspect.getargspec` is gone
in Python >= 3.11, so getfullargspec is used; behaviour is otherwise identical,
including the ValueError text on failure (:1272-1273).

`in_features` derivation from Data.feature_extractor follows code/train.py:392-397.
"""
import importlib
import inspect
import os
import sys

_CODE_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "code")

FEATURE_DIMS = {"retccl": 2048, "histoencoder": 384, "ctranspath": 784}   # code/train.py:392-397


def _ensure_models_on_path():
    if _CODE_DIR not in sys.path:
        sys.path.insert(0, _CODE_DIR)


def load_model(model_cfg, **other_args):
    """model_cfg: mapping with at least `name` (the YAML `Model:` section)."""
    _ensure_models_on_path()
    get = model_cfg.get if hasattr(model_cfg, "get") else (lambda k, d=None: getattr(model_cfg, k, d))
    name = get("name")
    camel_name = "".join(i.capitalize() for i in name.split("_")) if "_" in name else name
    try:
        Model = getattr(importlib.import_module(f"models.{name}"), camel_name)
    except Exception:
        raise ValueError("Invalid Module File Name or Invalid Class Name!")
    class_args = inspect.getfullargspec(Model.__init__).args[1:]
    keys = list(model_cfg.keys())
    args = {a: model_cfg[a] for a in class_args if a in keys}
    args.update(other_args)
    return Model(**args)


def model_cfg_from_yaml(cfg):
    """Applies train.py's overrides to a parsed YAML dict and returns the `Model` section."""
    model = dict(cfg.get("Model", {}))
    fe = (cfg.get("Data", {}) or {}).get("feature_extractor")
    if fe in FEATURE_DIMS:
        model["in_features"] = FEATURE_DIMS[fe]
    return model


def load_reference_checkpoint(model, checkpoint, strict=True, map_location="cpu"):
    """Loads a checkpoint written by the reference's Lightning run into `model` (a TransMIL of this package).

    The reference saves `ModelInterface.state_dict()`, whose TransMIL parameters carry the attribute prefix `model.`, and
    restores them by replacing that prefix in every key (code/visualize_mil.py:165-172).  Same here; `checkpoint` is the
    dict `torch.load` returns (or a path to it), with or without the outer `state_dict` entry.  Keys that do not belong
    to the model (e.g. loss buffers of the LightningModule) are dropped when `strict` is False and raise, as the
    reference's strict load does, when it is True."""
    import torch
    if isinstance(checkpoint, (str, os.PathLike)):
        checkpoint = torch.load(checkpoint, map_location=map_location)
    weights = checkpoint.get("state_dict", checkpoint)
    renamed = {}
    for key, value in weights.items():
        renamed[key.replace("model.", "")] = value
    if not strict:
        own = set(model.state_dict().keys())
        renamed = {k: v for k, v in renamed.items() if k in own}
    return model.load_state_dict(renamed, strict=strict)


def step(model, input):
    """ModelInterface.step (code/models/model_interface.py:319-331): logits, Y_prob, Y_hat of one batch of bags."""
    import torch
    logits = model(input.float().contiguous())
    return logits, torch.softmax(logits, dim=1), torch.argmax(logits, dim=1)


def training_step(model, batch, loss_fn=None):
    """ModelInterface.training_step (code/models/model_interface.py:333-383) without the Lightning bookkeeping: `batch` is
    (input [B,n,D], label [B], anything), the loss is taken against the one-hot label (:346-347; default CrossEntropyLoss with
    probability targets, the reference's `create_loss('CrossEntropyLoss')`).  Returns the loss (1-element tensor, :348-349);
    `.backward()` on it runs the library's backward kernels (training.TransMILFunction)."""
    import torch
    input, label = batch[0], batch[1]
    logits, Y_prob, Y_hat = step(model, input)
    one_hot = torch.nn.functional.one_hot(label.to(logits.device), num_classes=model.n_classes).float()
    loss = (loss_fn or torch.nn.functional.cross_entropy)(logits, one_hot)
    if loss.ndim == 0:
        loss = loss.unsqueeze(0)
    return loss, Y_prob, Y_hat


def cross_entropy_torch(x, y):
    """code/utils/utils.py:242-247: mean over the batch of -log softmax(x_i)[y_i], as a detached scalar (the reference builds
    it with torch.tensor(...), so it carries no graph)."""
    import torch
    y = torch.as_tensor(y, device=x.device).long().reshape(-1)
    return -torch.log_softmax(x.detach().float(), dim=1).gather(1, y[:, None]).sum() / y.shape[0]


def validation_step(model, batch):
    """ModelInterface.validation_step (code/models/model_interface.py:432-470) without the Lightning bookkeeping: `batch` is
    (input, label, (wsi_name, coords, patient)); returns the dict the reference appends to `validation_step_outputs` (:470):
    detached logits / Y_prob / Y_hat, `label.int()`, slide name, patient and the `cross_entropy_torch` loss (:446).  Runs the
    inference path (no autograd) whatever mode the module is in, as Lightning's validation loop does."""
    import torch
    input, label, (wsi_name, _, patient) = batch
    was_training = bool(getattr(model, "training", False))
    if was_training:
        model.eval()            # Lightning's validation loop puts the module in eval(): Dropout off
    try:
        with torch.no_grad():
            logits, Y_prob, Y_hat = step(model, input)
    finally:
        if was_training:
            model.train()
    logits, Y_prob, Y_hat = logits.detach(), Y_prob.detach(), Y_hat.detach()
    loss = cross_entropy_torch(logits, label)
    return {"logits": logits, "Y_prob": Y_prob, "Y_hat": Y_hat, "label": torch.as_tensor(label).int(), "name": wsi_name,
            "patient": patient, "loss": loss}
