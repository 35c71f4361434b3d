from .mnist_cnn import (build_model_a, build_model_b, build_and_compile_cnn_model,  # noqa: F401
                        build_and_compile_model_b)
