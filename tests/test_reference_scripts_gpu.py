"""GPU test: the reference's OWN files run through the B200 path.

tools/run_reference_script.py executes the unmodified main.py and main_dist.py (from the copy staged under
oracle/_ref/reference by oracle/stage_reference.py — the GPU box has no /root/reference) for one short epoch on a
synthetic data set, once with accelerate() under the script's module and once plain (the reference's own torch /
cuDNN fp32 path = the oracle here). Same seed, so same initial weights, same shuffling and the same crop / flip
draws in both runs; compared: every criterion() value of train() and test(), and the checkpoint file the script
wrote (`module.`-prefixed state_dict, main.py:138-147) re-loaded into a plain fp32 model.
Tolerances: bf16 storage against fp32 on a loss around 2.3 — first step 3e-2, later steps (the two runs' weights
drift apart by the storage noise) 0.15.
"""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run(tmp_path, tag, script, plain, script_args, model="resnet.ResNet18", extra=()):
    from oracle.stage_reference import reference_dir
    if reference_dir() is None:
        pytest.fail("reference files are not staged: run `python oracle/stage_reference.py` (or __graft_entry__.build()) "
                    "on a machine that has /root/reference before shipping the snapshot")
    out = str(tmp_path / f"{tag}.json")
    cmd = [sys.executable, os.path.join(ROOT, "tools", "run_reference_script.py"), "--script", script, "--model", model,
           "--out", out] + (["--plain"] if plain else []) + list(extra) + ["--"] + script_args
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="0")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    return json.load(open(out))


def compare(acc, ref, steps):
    assert acc["accelerated"] and acc["kernels_used"] and not ref["kernels_used"]
    assert acc["sgd_steps"] == ref["sgd_steps"] == steps
    assert len(acc["train_losses"]) == len(ref["train_losses"]) == steps
    assert abs(acc["train_losses"][0] - ref["train_losses"][0]) < 3e-2, (acc["train_losses"], ref["train_losses"])
    for a, b in zip(acc["train_losses"], ref["train_losses"]):
        assert abs(a - b) < 0.15, (acc["train_losses"], ref["train_losses"])
    assert len(acc["test_losses"]) == len(ref["test_losses"]) > 0
    for a, b in zip(acc["test_losses"], ref["test_losses"]):
        assert abs(a - b) < 0.15, (acc["test_losses"], ref["test_losses"])
    ck = acc["ckpt"]
    assert ck["top_level_keys"] == ["acc", "epoch", "net"] and ck["module_prefix"]
    assert ck["n_entries"] == ref["ckpt"]["n_entries"] >= 122
    assert "All keys matched" in ck["load_result"] and ck["epoch"] == 0
    mean_test = sum(acc["test_losses"]) / len(acc["test_losses"])
    assert abs(ck["plain_eval_loss"] - mean_test) < 0.1          # the saved weights in a plain fp32 model


@pytest.mark.parametrize("model", ["dla_simple.SimpleDLA", "resnet.ResNet18", "senet.SENet18", "googlenet.GoogLeNet",
                                   "resnext.ResNeXt29_2x64d"])
def test_main_py_unmodified(tmp_path, model):
    """main.py: 428 samples = 3 batches of 128 + a ragged 44, then test() on 2 x 100 and the checkpoint.
    SimpleDLA is the network main.py:71 itself constructs; ResNet-18 is the benchmarked one (main.py:58); SENet18,
    GoogLeNet and ResNeXt29_2x64d are three more of the constructors main.py:57-70 offers."""
    args = ["--lr", "0.02"]
    acc = run(tmp_path, "main_acc", "main.py", False, args, model=model)
    ref = run(tmp_path, "main_ref", "main.py", True, args, model=model)
    compare(acc, ref, 4)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(dict(accelerated=acc, plain=ref),
              open(os.path.join(ROOT, "gpurun_out", f"reference_main_py_{model.split('.')[-1]}.json"), "w"), indent=1)


def test_main_dist_py_unmodified_under_ddp(tmp_path):
    """main_dist.py --dist: NCCL process group, DistributedSampler, DistributedDataParallel(net, device_ids=[gpu])
    around the accelerated module (main_dist.py:73,110,141-144), with the arena-bucket all-reduce attached and with
    DDP's own Reducer doing the all-reduce."""
    def args(port, tag):
        out = str(tmp_path / f"out_{tag}")
        return ["--dist", "--world_size", "1", "--rank", "0", "--dist_url", f"tcp://127.0.0.1:{port}", "--batch_size",
                "128", "--epochs", "1", "--output_dir", out, "--workers", "2", "--lr", "0.02"]
    acc = run(tmp_path, "dist_acc", "main_dist.py", False, args(29541, "a"))
    red = run(tmp_path, "dist_red", "main_dist.py", False, args(29542, "b"), extra=["--no-attach"])
    ref = run(tmp_path, "dist_ref", "main_dist.py", True, args(29543, "c"))
    assert acc["ddp_attached"] and not red["ddp_attached"]
    compare(acc, ref, 4)
    compare(red, ref, 4)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(dict(attached=acc, reducer=red, plain=ref),
              open(os.path.join(ROOT, "gpurun_out", "reference_main_dist_py.json"), "w"), indent=1)


def test_main_dist_py_amp_flag(tmp_path):
    """main_dist.py --amp (main_dist.py:68-69,179-186): the script wraps forward and loss in torch.cuda.amp.autocast() and
    drives the optimizer through GradScaler (scale(loss).backward(), step, update). With accelerate() underneath, the
    protocol is honoured as it stands — the plan is one autograd node autocast does not touch, the scaled loss
    gradient flows through the bf16-storage backward (bf16 has fp32's exponent range, so the 2^16 scale cannot
    overflow it) and GradScaler unscales the arena-backed .grad tensors — while the arithmetic stays this library's
    bf16 storage / fp32 accumulation rather than fp16. Checked against the same script WITHOUT --amp on plain torch
    fp32: same losses within the bf16 tolerance, 4 optimizer steps taken (no step skipped for inf / nan)."""
    def args(port, tag, amp):
        out = str(tmp_path / f"out_{tag}")
        return ["--dist", "--world_size", "1", "--rank", "0", "--dist_url", f"tcp://127.0.0.1:{port}", "--batch_size",
                "128", "--epochs", "1", "--output_dir", out, "--workers", "2", "--lr", "0.02"] + (["--amp"] if amp else [])
    acc = run(tmp_path, "amp_acc", "main_dist.py", False, args(29551, "a", True))
    ref = run(tmp_path, "amp_ref", "main_dist.py", True, args(29552, "b", False))
    compare(acc, ref, 4)
