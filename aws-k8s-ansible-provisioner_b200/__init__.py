"""aws-k8s-ansible-provisioner_b200 -- B200-native (sm_100a) model-server engine for the one hot path the
reference deploys: the Qwen3-0.6B forward pass inside the llm-d model-server pod
(reference: llm-d-deploy.yaml:118,176-184; llm-d-test.yaml:32-73).

The package holds the CUDA/C++ engine (csrc/, built in-tree into libb2engine.so and the
b2serve binary) and a thin ctypes mirror of its C ABI (include/b2_engine.h).  There is no CPU
fallback: importing works anywhere, but every compute entry point fails loudly without the
built library and a B200.

The directory name carries the reference's hyphens, so import it with
    importlib.import_module("aws-k8s-ansible-provisioner_b200")
(`__graft_entry__.py` and tests/conftest.py do exactly that).
"""
from .capi import LIB_PATH, load_library, B2Error  # noqa: F401
from .engine import Engine, EngineConfig, QWEN3_0_6B, KERNEL_CLASSES  # noqa: F401
