from .resnet import ResNet, resnet18, resnet50
from .workshop_net import Net, WorkshopNet


def get_model(model_type: str):
    """``--model-type`` dispatch of the reference (…smddp-gpu.py:30-52): 'resnet18' -> torchvision
    topology, anything else ('custom') -> the workshop ``Net``.  'resnet50' is the extra BASELINE config."""
    if model_type == "resnet18":
        return resnet18()
    if model_type == "resnet50":
        return resnet50()
    return WorkshopNet()


__all__ = ["ResNet", "resnet18", "resnet50", "Net", "WorkshopNet", "get_model"]
