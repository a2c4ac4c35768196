"""honk_b200 - B200-native batched keyword-spotting inference behind Honk's own API.

``honk_b200.manage_audio`` and ``honk_b200.model`` mirror the reference's
``utils.manage_audio`` / ``utils.model`` for the inference hot path;
``honk_b200.service`` mirrors ``service.TorchLabelService``.
"""
from .model import (ConfigType, SerializableModule, SpeechModel, SpeechResModel,  # noqa: F401
                    find_config, find_model)
from .manage_audio import AudioPreprocessor  # noqa: F401

__all__ = ["AudioPreprocessor", "ConfigType", "SerializableModule", "SpeechModel", "SpeechResModel",
           "find_config", "find_model"]
