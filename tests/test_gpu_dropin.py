"""GPU: the drop-in, proven from the reference side.  A subprocess loads the shim built under the
reference's library names (_C / _C_cache_ops / _moe_C) INSTEAD of the installed vllm wheel's extension and drives the
hot-path ops through the wheel's own `vllm/_custom_ops.py` wrappers (tests/dropin/drive_wheel_custom_ops.py); a second
one resolves the attention backend through the platform plugin's entry point object."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _wheel_present():
    import importlib.util
    return importlib.util.find_spec("vllm") is not None


def test_wheel_custom_ops_drive_the_dropin_shim():
    if not _wheel_present():
        pytest.skip("no vllm wheel in this environment")
    so = os.path.join(ROOT, "vllm_b200", "lib", "libvllm_b200_dropin.so")
    assert os.path.exists(so), "drop-in shim not built (run __graft_entry__.build())"
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "dropin", "drive_wheel_custom_ops.py")],
                       capture_output=True, text=True, timeout=900, cwd="/tmp")
    out = r.stdout + r.stderr
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("[dropin]") or ln.startswith("dropin:")]
    print("\n".join(lines))
    assert r.returncode == 0, out[-4000:]
    assert "dropin: " in r.stdout and " ops OK" in r.stdout


def test_platform_plugin_resolves_backend_on_gpu():
    if not _wheel_present():
        pytest.skip("no vllm wheel in this environment")
    code = (
        "import sys; sys.path.insert(0, %r)\n"
        "import vllm_b200_plugin as p\n"
        "qual = p.register()\n"
        "assert qual == 'vllm_b200_plugin.platform.B200Platform', qual\n"
        "from vllm_b200_plugin.platform import resolve_platform\n"
        "plat = resolve_platform()\n"
        "cls = plat.get_attn_backend_cls(None, None)\n"
        "assert cls == 'vllm_b200.attention.backend.B200PagedAttentionBackend', cls\n"
        "import importlib; m, c = cls.rsplit('.', 1); B = getattr(importlib.import_module(m), c)\n"
        "assert B.get_name() == 'B200_PAGED'\n"
        "plat.import_kernels(); import torch\n"
        "assert hasattr(torch.ops._C, 'paged_attention_v1') and hasattr(torch.ops._C_cache_ops, 'reshape_and_cache')\n"
        "print('plugin: OK', plat.device_name)\n" % ROOT)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600, cwd="/tmp")
    assert r.returncode == 0, (r.stdout + r.stderr)[-4000:]
    assert "plugin: OK" in r.stdout
