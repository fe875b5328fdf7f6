from nanofed_b200.core.types import ModelVersion

from .manager import ModelManager, make_json_serializable

__all__ = ["ModelManager", "ModelVersion", "make_json_serializable"]
