"""Model zoo.  ``create_model(arch)`` mirrors ``torchvision.models.__dict__[arch]()``
(/root/reference/distributed.py:132-137): the ResNet family is served by the
fused B200 implementation (same state_dict), every other torchvision name is
instantiated from torchvision and still trains under our parallel wrappers,
optimizer and loss kernels."""
import os

from .resnet import ResNet, BasicBlock, Bottleneck, build_resnet, native_archs


def create_model(arch, pretrained=False, num_classes=None):
    import torchvision.models as tvm
    use_native = arch in native_archs() and os.environ.get("PDT_MODEL_IMPL", "native") == "native"
    if use_native:
        model = build_resnet(arch, num_classes or 1000)
        if pretrained:   # needs the torchvision weight files (network), like the reference
            tv = tvm.__dict__[arch](pretrained=True)
            model.load_state_dict(tv.state_dict())
        return model
    kwargs = {} if num_classes is None else {"num_classes": num_classes}
    return tvm.__dict__[arch](pretrained=True, **kwargs) if pretrained else tvm.__dict__[arch](**kwargs)


__all__ = ["create_model", "ResNet", "BasicBlock", "Bottleneck", "build_resnet", "native_archs"]
