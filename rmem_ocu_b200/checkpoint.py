"""Checkpoint ingestion with the reference's ``load_network`` semantics (utils/checkpoint.py:75-104), so that a
released RMem checkpoint (e.g. ``aotplus_R50_DeAOTL_Temp_pe_Slot_4_ema_20000.pth``) loads into either the
reference's own model object or this package's parameter containers (rmem_ocu_b200/model.py carries the
reference's key names and shapes):

  * the weights sit under ``'state_dict'``, else under ``'model'``, else the file is the state dict itself;
  * a tensor whose key is in the model, has more than two dimensions and exactly ONE input channel fewer than the
    model's (``shape[1] == model.shape[1] - 1``, same ``shape[0]``) fills the leading channels and leaves the model's
    last channel as initialised -- this is how an 11-class ``patch_wise_id_bank`` checkpoint enters a model built with
    the extra ignore channel (MODEL_IGNORE_TOKEN);
  * a key found in the model with an equal shape is taken;
  * a key with the DataParallel prefix ``module.`` is taken under its stripped name if that matches in name and
    shape (and silently dropped if it does not);
  * every other key is reported back as not used; model keys the checkpoint lacks keep their initial values.
"""
import torch


def select_state_dict(pretrained):
    """utils/checkpoint.py:79-84."""
    if "state_dict" in pretrained.keys():
        return pretrained["state_dict"]
    if "model" in pretrained.keys():
        return pretrained["model"]
    return pretrained


def merge_state_dict(model_dict, pretrained_dict):
    """The merge rule of load_network (utils/checkpoint.py:85-100) on plain dicts of tensors.  ``model_dict`` is
    updated in place (as the reference does through the live parameter tensors) and returned together with the list of
    checkpoint keys that were not used."""
    update, removed = {}, []
    for k, v in pretrained_dict.items():
        if k in model_dict and (len(v.shape) > 2 and v.shape[0] == model_dict[k].shape[0]
                                and v.shape[1] == (model_dict[k].shape[1] - 1)):
            with torch.no_grad():
                model_dict[k][:, :-1, :, :] = v
            continue
        if k in model_dict and v.shape == model_dict[k].shape:
            update[k] = v
        elif k[:7] == "module.":
            if k[7:] in model_dict and v.shape == model_dict[k[7:]].shape:
                update[k[7:]] = v
        else:
            removed.append(k)
    model_dict.update(update)
    return model_dict, removed


def load_network(net, pretrained_dir, gpu=-1):
    """utils/checkpoint.py:75-104: returns (net, keys of the checkpoint that were not used)."""
    device_map = "cpu" if gpu == -1 or not torch.cuda.is_available() else torch.device("cuda:" + str(gpu))
    pretrained = torch.load(pretrained_dir, map_location=device_map, weights_only=False)
    model_dict, removed = merge_state_dict(net.state_dict(), select_state_dict(pretrained))
    net.load_state_dict(model_dict)
    # engines cache the device-side repack of a model's weights on the model object: new weights invalidate it
    net.__dict__.pop("_rmem_prepared", None)
    if gpu != -1 and torch.cuda.is_available():
        net = net.cuda(gpu)
    return net, removed
