"""CPU: the platform plugin's entry point (group `vllm.platform_plugins`, setup.py) loads and answers the way
vLLM's plugin loader expects (/root/reference/vllm/plugins/__init__.py:29-70: `plugin.load()` -> callable ->
Optional[qualified class name])."""
import importlib.metadata as md
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_entry_point():
    src = open(os.path.join(ROOT, "setup.py")).read()
    m = re.search(r'"vllm\.platform_plugins":\s*\["(\w+)\s*=\s*([\w\.]+:\w+)"\]', src)
    assert m, "setup.py must declare the vllm.platform_plugins entry point"
    return md.EntryPoint(name=m.group(1), value=m.group(2), group="vllm.platform_plugins")


def test_entry_point_loads_and_gates_on_the_device(monkeypatch):
    sys.path.insert(0, ROOT)
    ep = _declared_entry_point()
    fn = ep.load()          # what load_plugins_by_group does
    assert callable(fn)
    monkeypatch.delenv("VLLM_B200_FORCE_PLUGIN", raising=False)
    import torch
    if not torch.cuda.is_available():
        assert fn() is None     # not a B200: vLLM keeps its built-in platform detection
    monkeypatch.setenv("VLLM_B200_FORCE_PLUGIN", "1")
    qual = fn()
    assert qual == "vllm_b200_plugin.platform.B200Platform"
    mod, cls = qual.rsplit(".", 1)
    assert os.path.exists(os.path.join(ROOT, *mod.split(".")) + ".py") and f"class {cls}(" in open(
        os.path.join(ROOT, *mod.split(".")) + ".py").read()
