from .spec import NetSpec, ConvBN, MaxPool, Head, enhanced_cnn_spec, resnet50_spec, get_spec  # noqa: F401
from .torch_net import SpecNet, EnhancedCNNModel, ResNet50, ResBlock, xavier_init_  # noqa: F401
