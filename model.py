"""Upstream module name kept: ``from model import UserModel, ServerUserModel``."""
from fedrec_b200.models.model import UserModel, ServerUserModel  # noqa: F401
