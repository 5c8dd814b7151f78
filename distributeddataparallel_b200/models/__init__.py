from .resnet import SimpleCNN, ResNet, BasicBlock, Bottleneck, resnet18, resnet34, resnet50


def build_gpt2(name: str = "gpt2-medium", seq_len: int = 1024):
    """GPT-2 small / medium / tiny (``models/gpt2.py``), imported lazily."""
    from .gpt2 import build_gpt2 as _build

    return _build(name, seq_len)


__all__ = ["SimpleCNN", "ResNet", "BasicBlock", "Bottleneck", "resnet18", "resnet34", "resnet50", "build_gpt2"]
