"""host/mrt_render.c, the headless plain-C driver that stands where the reference's windowed frame loop is
(-drawInMTKView:, Renderer.mm:587-638): what can be checked without a GPU — malformed arguments are refused before
anything is created (exit code 2, one line on stderr), host-side failures exit 1 with the library's message, and without a
CUDA device the driver FAILS with the library's "no CPU fallback" error instead of rendering something else."""
import pathlib
import subprocess

import pytest

ROOT = pathlib.Path(__file__).resolve().parents[1]
BIN = ROOT / "metal-renderer_b200" / "bin" / "mrt_render"


@pytest.fixture(scope="module", autouse=True)
def built_driver():
    if not BIN.exists():                                 # conftest builds the library; the driver is one more make target
        subprocess.check_call(["make", "-s", "-C", str(BIN.parents[1]), "bin/mrt_render"])


def run(*args):
    p = subprocess.run([str(BIN), *args], capture_output=True, text=True, timeout=120)
    return p.returncode, p.stdout, p.stderr


@pytest.mark.parametrize("args,message", [
    (["--optics", "rainbow"], "bad --optics"), (["--reduce", "mpi"], "bad --reduce"), (["--size", "12"], "bad --size"),
    (["--shard", "tiles"], "bad --shard"), (["--pipeline", "hybrid"], "bad --pipeline"), (["--frobnicate"], "unknown argument --frobnicate"),
    (["--devices", "0,1", "--tone"], "need a single device"), (["--devices", "0,1", "--checkpoint", "x.ckpt"], "need a single device"),
])
def test_malformed_arguments_exit_2_before_anything_is_created(args, message):
    rc, out, err = run(*args)
    assert rc == 2 and message in err and out == ""


def test_host_side_failures_exit_1_with_the_library_message(tmp_path):
    rc, out, err = run("--scene", "no-such-scene")
    assert rc == 1 and "unknown procedural scene 'no-such-scene'" in err and out == ""
    rc, out, err = run("--scene", str(tmp_path / "absent.obj"))
    assert rc == 1 and "cannot open" in err and out == ""


def test_without_a_gpu_the_driver_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("this machine has a CUDA device; the no-device path is what is tested")
    rc, out, err = run("--scene", "cornell", "--size", "32x32", "--frames", "1")
    assert rc == 1 and "no CUDA device" in err and "no CPU fallback" in err and out == ""
