"""Regenerate the fixtures that come from the reference's own header (needs /root/reference)."""
import pathlib
import subprocess

root = pathlib.Path(__file__).resolve().parents[2]
subprocess.check_call(["make", "-C", str(root / "oracle"), "_ref/ref_probe"])
probe = root / "oracle" / "_ref" / "ref_probe"
(root / "tests/golden/abi_layout.json").write_text(subprocess.check_output([str(probe)], text=True))
(root / "tests/golden/ref_host_kats.json").write_text(subprocess.check_output([str(probe), "k"], text=True))
print("wrote abi_layout.json, ref_host_kats.json")
