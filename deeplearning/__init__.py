"""Namespace package root: the B200 tensor backend for DeepLearning.scala lives in `deeplearning.scala_b200`."""
