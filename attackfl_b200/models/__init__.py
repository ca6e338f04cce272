"""Model registry.  The reference resolves models by class name with ``getattr(src.Model, name)``
(``server.py:139-140,169-170``; ``src/RpcClient.py:74-75``; ``src/Validation.py:25-26``); ``build``
keeps that contract, including the error text."""
from .icu import CNNModel, RNNModel, TransformerBlock, TransformerModel, BRANCHES, VITALS_DIM, LABS_DIM
from .har import PositionalEncoding, TransformerClassifier, HAR_FEATURES, HAR_CLASSES
from .hyper import HyperNetwork, CNNHyper, FusedHyperNet

_REGISTRY = {
    "CNNModel": CNNModel,
    "RNNModel": RNNModel,
    "TransformerModel": TransformerModel,
    "TransformerClassifier": TransformerClassifier,
}


def build(model_name: str, **kwargs):
    if model_name not in _REGISTRY:
        raise ValueError(f"Model name '{model_name}' is not valid.")
    return _REGISTRY[model_name](**kwargs)


def names():
    return list(_REGISTRY)


__all__ = ["CNNModel", "RNNModel", "TransformerBlock", "TransformerModel", "PositionalEncoding",
           "TransformerClassifier", "HyperNetwork", "CNNHyper", "FusedHyperNet", "build", "names",
           "BRANCHES", "VITALS_DIM", "LABS_DIM", "HAR_FEATURES", "HAR_CLASSES"]
