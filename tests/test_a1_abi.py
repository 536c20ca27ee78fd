"""The C-ABI library loads and exports every symbol include/oflow_b200.h declares
(no compute calls: runs without a GPU)."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    names = set()
    inc = os.path.join(ROOT, "include")
    for f in os.listdir(inc):
        if f.endswith(".h"):
            txt = open(os.path.join(inc, f)).read()
            txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
            names |= set(re.findall(r"\b(of_[a-z0-9_]+)\s*\(", txt))
    return names


def test_header_symbols_exported(lib):
    from optical_flow_b200 import _lib
    declared = _declared_symbols()
    assert len(declared) >= 25
    for name in sorted(declared):
        assert hasattr(lib, name), "missing export " + name
    # every declared symbol has a ctypes prototype and vice versa
    assert declared == set(_lib.PROTOTYPES.keys())


def test_version_and_error_string(lib):
    assert lib.of_version() >= 100
    assert isinstance(lib.of_last_error(), bytes)


def test_argument_errors_without_gpu(lib):
    # shape/argument validation happens before any CUDA call
    from optical_flow_b200 import _lib
    rc = lib.of_warp_features_fwd(None, None, None, None, 1, 4, 4, 4, 0, 0, None)
    assert rc == -1
    assert b"null" in lib.of_last_error()
    rc = lib.of_cost_volume_fwd(1, 1, 1, 1, 4, 4, 6, 3, 49, 0, None)   # C % 4 != 0
    assert rc == -2
    try:
        _lib.check(rc, "of_cost_volume_fwd")
        assert False
    except AssertionError as e:
        assert "C % 4" in str(e)


def test_no_product_import_of_oracle():
    """The product package must never import the oracle (it is test infrastructure)."""
    pkg = os.path.join(ROOT, "optical_flow_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                txt = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", txt, flags=re.M), os.path.join(dirpath, f)


def test_regularized_parameter_names_follow_the_reference():
    """Host logic of the opt-in L2 term: the reference attaches l2(1e-4) to kernel AND bias of the stem (model.py:12-13)
    and of each residual block's conv1 / conv2 (resnet/models.py:17-23) - 13 convs, not the shortcuts, BN or decoder."""
    from optical_flow_b200.params import decoder_param_specs, encoder_param_specs
    from optical_flow_b200.train import regularized_parameters
    specs = dict(encoder_param_specs())
    specs.update(decoder_param_specs())
    names = regularized_parameters(specs)
    assert len(names) == 26 and len(set(names)) == 26
    assert {"conv1/kernel", "conv1/bias", "layer4_block_2_conv2/kernel", "layer3_block_1_conv1/bias"} <= set(names)
    assert not any("shortcut" in n or "_bn" in n or n.startswith(("flow", "layer1_bn")) for n in names)
