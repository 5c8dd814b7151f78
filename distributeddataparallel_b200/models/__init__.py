from .resnet import SimpleCNN, ResNet, BasicBlock, Bottleneck, resnet18, resnet34, resnet50


def gpt2(name: str = "gpt2-medium", seq_len: int = 1024):
    from .gpt2 import build_gpt2

    return build_gpt2(name, seq_len)


__all__ = ["SimpleCNN", "ResNet", "BasicBlock", "Bottleneck", "resnet18", "resnet34", "resnet50", "gpt2"]
