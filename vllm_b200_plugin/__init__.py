"""vLLM platform plugin that puts this repo's B200 kernels behind an unmodified vLLM.

Registration (setup.py at the repo root):
    entry_points = {"vllm.platform_plugins": ["b200 = vllm_b200_plugin:register"]}
vLLM discovers the group in `load_plugins_by_group("vllm.platform_plugins")`
(/root/reference/vllm/platforms/__init__.py:190-218, /root/reference/vllm/plugins/__init__.py:29-70), calls the
function, and -- when it returns a qualified class name -- uses that class as `current_platform`.  Model:
/root/reference/tests/plugins/vllm_add_dummy_platform (setup.py entry point + `dummy_platform_plugin()`).

`register()` activates only on a Blackwell data-centre GPU (compute capability 10.x); anywhere else it returns None
and vLLM carries on with its built-in platform detection.  VLLM_B200_FORCE_PLUGIN=1 forces it (tests)."""
import os
from typing import Optional

PLATFORM_CLS = "vllm_b200_plugin.platform.B200Platform"


def register() -> Optional[str]:
    if os.environ.get("VLLM_B200_FORCE_PLUGIN") == "1":
        return PLATFORM_CLS
    try:
        import torch
        if torch.cuda.is_available() and torch.cuda.get_device_capability(0)[0] == 10:
            return PLATFORM_CLS
    except Exception:  # noqa: BLE001  (no driver, no device: not our platform)
        pass
    return None
