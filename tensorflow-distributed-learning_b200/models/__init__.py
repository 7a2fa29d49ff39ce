"""Model zoo: the reference CNN (tf_dist_example.py:39-53) plus the larger configs named in BASELINE.json."""
from .bert import BertMLM, build_and_compile_bert_mlm
from .mnist_cnn import build_and_compile_cnn_model, build_cnn_model
from .resnet50 import ResNet50, build_and_compile_resnet50

__all__ = ["build_and_compile_cnn_model", "build_cnn_model", "ResNet50", "build_and_compile_resnet50", "BertMLM",
           "build_and_compile_bert_mlm"]
