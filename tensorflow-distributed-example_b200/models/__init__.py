from .mnist_cnn import (build_model_a, build_model_b, build_and_compile_cnn_model,  # noqa: F401
                        build_and_compile_model_b)
from .resnet import Bottleneck, build_resnet, build_resnet50, build_and_compile_resnet50  # noqa: F401,E402
