"""tools/validate_quantized.py keeps the reference's click flags and its 4-line result file.

CPU part: every option the reference declares in run_method_options / approx_options
(utils/click_options.py:483-542 -- tests/golden/click_flags.json, extracted from the reference source by
tests/golden/make_click_flags.py) exists in our command with the same spelling (including the --no- twin), the same
default and the same kind; every flag scripts/image_net.sh and scripts/generated_scripts.py pass is accepted.
GPU part: one tiny validate-quantized run writes approx_output/<arch>/E{e}M{m}D{d}/D{d}_<timestamp>.txt whose four
lines have the reference's keys, order and literal format (image_net.py:185-202; the reference's own
approx_output/vit_quantized/E3M4D3/*.txt is the model)."""
import ast
import json
import os
import re
import sys

import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, os.path.join(ROOT, "tools"))
FLAGS = json.load(open(os.path.join(ROOT, "tests", "golden", "click_flags.json")))


def _command():
    import validate_quantized as vq
    return vq.validate_quantized


def test_reference_flag_set_is_accepted_unchanged():
    cmd = _command()
    by_opt = {}
    for p in cmd.params:
        for o in list(p.opts) + list(p.secondary_opts):
            by_opt[o] = p
    for group in ("run_method_options", "approx_options"):
        for ref in FLAGS[group]:
            on, _, off = ref["decl"].partition("/")
            assert on in by_opt, f"{group}: {on} missing"
            p = by_opt[on]
            if off:
                assert off in p.secondary_opts, f"{on}: --no- twin {off} missing"
                assert p.is_flag
            if "default" in ref:
                assert p.default == ref["default"], (on, p.default, ref["default"])
            if ref.get("type") == "int":
                assert p.type.name == "integer", on
    missing = [f for f in FLAGS["script_flags"] if f not in by_opt]
    assert not missing, missing


def test_fixture_matches_the_reference_source():
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import make_click_flags as mk
    import ref_shim
    if not ref_shim.available():
        pytest.skip("reference not present")
    tree = ast.parse(open(os.path.join(ref_shim.REF_ROOT, "utils", "click_options.py")).read())
    for group in ("run_method_options", "approx_options"):
        assert mk.options_of(tree, group) == FLAGS[group]


@pytest.mark.gpu
def test_result_file_format(tmp_path):
    from click.testing import CliRunner
    import validate_quantized as vq
    out_dir = str(tmp_path / "approx_output")
    args = ("validate-quantized --architecture resnet18_quantized_approx --batch-size 2 --seed 10 --n-bits 8 --cuda "
            "--qmethod fp_quantizer --per-channel --fp8-mantissa-bits=3 --fp8-set-maxval "
            "--weight-quant-method=current_minmax --act-quant-method=allminmax --num-est-batches=1 --quantize-input "
            "--approx_flag --res-quantizer-flag --no-original-quantize-res --expo-width 4 --mant-width 3 "
            "--dnsmp-factor 3 --withComp --with_approx --with_s2nn2s_opt --quant_btw_mult_accu --num-eval-batches 1 "
            f"--approx-output-dir {out_dir}").split()
    res = CliRunner().invoke(vq.cli, args, catch_exceptions=False)
    assert res.exit_code == 0, res.output
    d = os.path.join(out_dir, "resnet18_quantized_approx", "E4M3D3")
    files = os.listdir(d)
    assert len(files) == 1 and re.fullmatch(r"D3_\d{4}-\d{2}-\d{2}_\d{2}-\d{2}-\d{2}\.txt", files[0]), files
    lines = open(os.path.join(d, files[0])).read().splitlines()
    assert [ln.split(":", 1)[0] for ln in lines] == ["evaluate_param", "run_method", "approx_params", "final_metrics"]
    assert lines[0] == "evaluate_param: mini_test"
    run_method = ast.literal_eval(lines[1].split(": ", 1)[1])
    approx_params = ast.literal_eval(lines[2].split(": ", 1)[1])
    metrics = ast.literal_eval(lines[3].split(": ", 1)[1])
    assert list(run_method) == ["approx_flag", "quantize_after_mult_and_add", "res_quantizer_flag", "original_quantize_res"]
    assert run_method == dict(approx_flag=True, quantize_after_mult_and_add=False, res_quantizer_flag=True,
                              original_quantize_res=False)
    assert list(approx_params) == ["expo_width", "mant_width", "dnsmp_factor", "withComp", "with_approx",
                                   "with_s2nn2s_opt", "sim_hw_add_OFUF", "with_OF_opt", "with_UF_opt", "golden_clip_OF",
                                   "quant_btw_mult_accu", "debug_mode", "self_check_mode"]
    assert approx_params["expo_width"] == 4 and approx_params["withComp"] is True
    assert list(metrics) == ["top_1_accuracy", "top_5_accuracy", "loss"]
    assert 0.0 <= metrics["top_1_accuracy"] <= 1.0 and metrics["loss"] >= 0.0
