"""haet_b200 -- B200-native (sm_100a) implementation of the HAET per-layer hot path.

Public surface (mirrors the reference's ``models.PhysicsAttention`` and ``balltree`` modules):
    Physics_Attention_Irregular_Mesh, Physics_Attention_Structured_Mesh_2D, Physics_Attention_Structured_Mesh_3D
    ErwinTransformer, build_balltree_with_rotations, build_balltree, patch_model, HostStream (host-resident meshes),
    block_mlp / Physics_Attention_*.forward_block (the two halves of TransolverErwinBlock.forward, fused)
    HAETransolver_Irregular_Mesh / _Structured_Mesh_2D / _Structured_Mesh_3D (the reference's Model constructors;
    forward(x, fx, T) = fused prologue + fused blocks + fused head)
All arithmetic runs in libhaet_b200.so (hand-written CUDA behind the C-ABI of include/haet_b200.h).
"""
from ._lib import HaetError, lib  # noqa: F401
from .balltree import build_balltree, build_balltree_with_rotations  # noqa: F401
from .erwin import ErwinTransformer  # noqa: F401
from .host_stream import HostStream, bind_host_to_gpu_numa  # noqa: F401
from .physics_attention import (Physics_Attention_Irregular_Mesh, Physics_Attention_Structured_Mesh_2D,  # noqa: F401
                                Physics_Attention_Structured_Mesh_3D, block_mlp, from_reference, patch_model)
from .models import (HAETransolver_Irregular_Mesh, HAETransolver_Structured_Mesh_2D,  # noqa: F401
                     HAETransolver_Structured_Mesh_3D, fused_model_forward, head_ln_linear, model_prologue)

__all__ = ["Physics_Attention_Irregular_Mesh", "Physics_Attention_Structured_Mesh_2D",
           "Physics_Attention_Structured_Mesh_3D", "ErwinTransformer", "build_balltree_with_rotations",
           "build_balltree", "patch_model", "block_mlp", "from_reference", "HostStream", "bind_host_to_gpu_numa", "HaetError", "lib",
           "HAETransolver_Irregular_Mesh", "HAETransolver_Structured_Mesh_2D", "HAETransolver_Structured_Mesh_3D",
           "fused_model_forward", "head_ln_linear", "model_prologue"]
