#!/usr/bin/env python
"""`image_net.py validate-quantized` on a synthetic ImageNet-shaped batch.

Takes the click flags of the reference's scripts unchanged (scripts/image_net.sh:23-62,
scripts/generated_scripts.py:20-58; option definitions utils/click_options.py:23-103, 320-542), builds the same
three dicts (quant_params_dict :544-577, run_method_dict :579-587, approx_params_dict :589-606), runs
estimate_ranges -> calibration batch(es) -> fix_ranges -> the 10-batch "mini_test" evaluation
(image_net.py:100-112, :171-179) and writes the reference's 4-line result file
approx_output/<arch>/E{e}M{m}D{d}/D{d}_<timestamp>.txt (image_net.py:185-202), plus throughput lines.

There is no ImageNet and no pretrained checkpoint on this box: images are seeded `randn`, labels are the
FP32 model's own argmax (so top-1 measures agreement with the unquantized network), weights are seeded
random.  `--images-dir` / `--model-dir` are accepted and ignored.  Multi-GPU: launch with torchrun; the eval
batches are sharded by rows and the logits gathered (fp8_quantization_b200/eval_harness.py).

  python tools/validate_quantized.py validate-quantized --architecture resnet18_quantized_approx \
      --batch-size 16 --seed 10 --n-bits 8 --cuda --qmethod fp_quantizer --per-channel --fp8-mantissa-bits=3 \
      --fp8-set-maxval --weight-quant-method=current_minmax --act-quant-method=allminmax --num-est-batches=1 \
      --quantize-input --approx_flag --res-quantizer-flag --no-original-quantize-res --expo-width 4 \
      --mant-width 3 --dnsmp-factor 3 --withComp --with_approx --with_s2nn2s_opt --quant_btw_mult_accu \
      --approx-output-dir approx_output
"""
import datetime
import json
import os
import sys
import time

import click

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

ARCHS = ["resnet18_quantized_approx", "resnet18_quantized", "mobilenet_v2_quantized_approx", "mobilenet_v2_quantized",
         "vit_quantized_approx", "vit_quantized"]


def flag(name, default=False):
    return click.option(f"--{name}/--no-{name}", is_flag=True, default=default)


@click.group()
def cli():
    pass


@cli.command("validate-quantized")
@click.option("--images-dir", default=None, help="accepted and ignored: synthetic images")
@click.option("--model-dir", default=None, help="accepted and ignored: seeded random weights")
@click.option("--architecture", type=click.Choice(ARCHS), required=True)
@click.option("--batch-size", type=int, default=16)
@click.option("--seed", type=int, default=10)
@click.option("--n-bits", type=int, default=8)
@click.option("--n-bits-act", type=int, default=None)
@flag("cuda", True)
@click.option("--load-type", default="fp32")
@click.option("--quant-setup", default="all")
@click.option("--qmethod", default="fp_quantizer")
@click.option("--qmethod-act", default=None)
@flag("per-channel")
@click.option("--fp8-mantissa-bits", type=int, default=4)
@flag("fp8-set-maxval")
@flag("fp8-mse-include-mantissa-bits", True)
@click.option("--weight-quant-method", default="current_minmax")
@click.option("--act-quant-method", default="allminmax")
@click.option("--num-est-batches", type=int, default=1)
@flag("quantize-input")
@flag("approx_flag")
@flag("quantize-after-mult-and-add")
@flag("res-quantizer-flag")
@flag("original-quantize-res")
@click.option("--expo-width", type=int, default=3)
@click.option("--mant-width", type=int, default=4)
@click.option("--dnsmp-factor", type=int, default=3)
@flag("withComp")
@flag("with_approx")
@flag("with_s2nn2s_opt")
@flag("sim_hw_add_OFUF")
@flag("with_OF_opt")
@flag("with_UF_opt")
@flag("golden-clip-OF")
@flag("quant_btw_mult_accu", True)
@flag("debug-mode")
@flag("self-check-mode")
@click.option("--approx-output-dir", default=None)
@click.option("--num-eval-batches", type=int, default=10, help='the reference\'s "mini_test" evaluates 10 batches')
def validate_quantized(**kw):
    import torch
    import torch.distributed as dist
    from fp8_quantization_b200 import eval_harness as eh, model_wrap as mw, ops, vit

    world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not kw["cuda"] or not torch.cuda.is_available():
        raise click.UsageError("fp8_quantization_b200 has no CPU path: run with --cuda on a B200")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    # click lower-cases option names (utils/click_options.py:527-533); re-cased as approx_params_dict does
    approx_params = dict(expo_width=kw["expo_width"], mant_width=kw["mant_width"], dnsmp_factor=kw["dnsmp_factor"],
                         withComp=kw["withcomp"], with_approx=kw["with_approx"], with_s2nn2s_opt=kw["with_s2nn2s_opt"],
                         sim_hw_add_OFUF=kw["sim_hw_add_ofuf"], with_OF_opt=kw["with_of_opt"],
                         with_UF_opt=kw["with_uf_opt"], golden_clip_OF=kw["golden_clip_of"],
                         quant_btw_mult_accu=kw["quant_btw_mult_accu"], debug_mode=kw["debug_mode"],
                         self_check_mode=kw["self_check_mode"])
    run_method = dict(approx_flag=kw["approx_flag"], quantize_after_mult_and_add=kw["quantize_after_mult_and_add"],
                      res_quantizer_flag=kw["res_quantizer_flag"], original_quantize_res=kw["original_quantize_res"])
    qparams = dict(method=kw["qmethod"], act_method=kw["qmethod_act"] or kw["qmethod"], n_bits=kw["n_bits"],
                   n_bits_act=kw["n_bits_act"] or kw["n_bits"], per_channel_weights=kw["per_channel"],
                   weight_range_method=kw["weight_quant_method"], act_range_method=kw["act_quant_method"],
                   quantize_input=bool(kw["quant_setup"] == "LSQ_paper" or kw["quantize_input"]),
                   fp8_kwargs=dict(mantissa_bits=kw["fp8_mantissa_bits"], set_maxval=kw["fp8_set_maxval"],
                                   mse_include_mantissa_bits=kw["fp8_mse_include_mantissa_bits"]),
                   custom_approx_params=approx_params, run_method=run_method)
    arch = kw["architecture"]
    torch.manual_seed(kw["seed"])
    if arch.startswith("resnet18"):
        from torchvision.models import resnet18
        fp_model = resnet18(weights=None)
        model = mw.QuantizedResNet(fp_model, **qparams)
    elif arch.startswith("mobilenet_v2"):
        fp_model = mw.MobileNetV2()
        model = mw.QuantizedMobileNetV2(fp_model, **qparams)
    else:
        fp_model = vit.VisionTransformerForImageClassification()
        model = vit.vit_quantized_approx(fp_model, **qparams)
    fp_model, model = fp_model.to(dev).eval(), model.to(dev).eval()
    bs = kw["batch_size"]
    gen = torch.Generator().manual_seed(kw["seed"])
    batch = lambda: torch.randn(bs, 3, 224, 224, generator=gen).to(dev)

    model.quantized()
    model.estimate_ranges()                                   # image_net.py:100-112
    with torch.inference_mode():
        for _ in range(kw["num_est_batches"]):
            model(batch())
    model.fix_ranges()
    top1 = top5 = n = 0
    loss = 0.0
    l0 = ops.kernel_launches()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    with torch.inference_mode():
        for _ in range(kw["num_eval_batches"]):               # "mini_test": 10 batches (image_net.py:171-179)
            x = batch()
            labels = fp_model(x).argmax(1)
            logits, m = eh.evaluate(model, x, labels)
            top1 += m["top1"] * bs
            top5 += m["top5"] * bs
            loss += float(torch.nn.functional.cross_entropy(logits, labels, reduction="sum"))
            n += bs
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    final_metrics = {"top_1_accuracy": top1 / n, "top_5_accuracy": top5 / n, "loss": loss / n}
    evaluate_param = "mini_test"
    if rank == 0:
        print(final_metrics)
        print(f"run_method: {run_method}")
        print(f"approx_params: {approx_params}")
        print(f"evaluate_param: {evaluate_param}")
        print(json.dumps({"images_per_s": round(n / dt, 1), "n_gpus": world, "images": n,
                          "kernel_launches": ops.kernel_launches() - l0, "labels": "argmax of the FP32 model",
                          "data": "synthetic"}))
        if kw["approx_output_dir"] is not None:               # image_net.py:185-202, same four lines
            out_dir = os.path.join(kw["approx_output_dir"], arch,
                                   "E{}M{}D{}".format(kw["expo_width"], kw["mant_width"], kw["dnsmp_factor"]))
            os.makedirs(out_dir, exist_ok=True)
            stamp = datetime.datetime.now().strftime("%Y-%m-%d_%H-%M-%S")
            path = out_dir + f"/D{kw['dnsmp_factor']}_{stamp}.txt"
            with open(path, "w") as f:
                f.write(f"evaluate_param: {evaluate_param}\n")
                f.write(f"run_method: {run_method}\n")
                f.write(f"approx_params: {approx_params}\n")
                f.write(f"final_metrics: {final_metrics}\n")
            print("wrote", path)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    cli()
