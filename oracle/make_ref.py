"""Recipe for oracle/_ref/: the reference's OWN implementation of the hot path, staged so that it can travel to the
GPU box (where /root/reference does not exist) and be timed there as the CPU baseline (`cpu_baseline.kind =
"reference"`, bench.py `--impl reference`) and cross-checked against the restatement in field_oracle.py.

The reference is pure Python: its per-ray path is the self-contained package /root/reference/models/
(tensoRF.py, tensorBase.py, mlp.py, sh.py — torch + numpy only).  "Building" it means
copying those files, unmodified, from where they lie into oracle/_ref/models/ (git-ignored: reference sources never
enter this repository's history).  Run by __graft_entry__.build() whenever /root/reference is present.

    python oracle/make_ref.py [/root/reference]
"""
import filecmp
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
FILES = ("__init__.py", "tensoRF.py", "tensorBase.py", "mlp.py", "sh.py")


def make(reference_root: str = "/root/reference") -> bool:
    src = os.path.join(reference_root, "models")
    if not os.path.isdir(src):
        return False
    dst = os.path.join(HERE, "_ref", "models")
    os.makedirs(dst, exist_ok=True)
    for name in FILES:
        s, d = os.path.join(src, name), os.path.join(dst, name)
        if not (os.path.exists(d) and filecmp.cmp(s, d, shallow=False)):
            shutil.copyfile(s, d)
            os.chmod(d, 0o644)
    return True


if __name__ == "__main__":
    ok = make(sys.argv[1] if len(sys.argv) > 1 else "/root/reference")
    print("oracle/_ref/models staged" if ok else "reference not found: nothing staged")
