from .bert import BertBase, BertLayer
from .layers import ConvBNAct, FusedLinear, Reshape, Reshape1
from .mobilenetv2 import Block, Block_nobn, MobileNetV2, MobileNetV2_nobn
from .resnet import ResNet, ResNet18, ResNet50
from .vgg import VGG, VGG16

MODEL_ZOO = {
    "mobilenetv2": MobileNetV2,
    "mobilenetv2_nobn": MobileNetV2_nobn,
    "resnet18": ResNet18,
    "resnet50": ResNet50,
    "vgg16": VGG16,
    "bert_base": BertBase,
}


def build_model(name: str, **kw):
    try:
        return MODEL_ZOO[name.lower()](**kw)
    except KeyError:
        raise ValueError("unknown model %r (have %s)" % (name, sorted(MODEL_ZOO)))
