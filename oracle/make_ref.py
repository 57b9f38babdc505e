"""TEST INFRASTRUCTURE ONLY -- packs the reference's own `model/` package, UNMODIFIED, into the
git-ignored `oracle/_ref/stemgnn_reference_model.zip` so that it travels to the GPU box (which has no
/root/reference) next to the built libraries.  Nothing is copied into the repository's history: the
archive is listed in .gitignore (`oracle/_ref/`), this recipe is what is committed.

What it is used for (oracle/ref_loader.py):
  * `bench.py --impl reference` and the `cpu_baseline` leg time the reference's OWN VectorQuantize
    (model/vq.py, torch + einops only) and its own Encoder (model/encoder.py) -- the latter on top of
    oracle/pyg_shim, because torch_geometric / torch_scatter are not installable here;
  * tests/test_gpu_reference_wrappers.py builds the reference's PretrainModel / TaskModel
    (model/pt_model.py, model/ft_model.py) over the drop-in modules.

    python oracle/make_ref.py            (a no-op with a message when /root/reference is absent)
"""
import hashlib
import os
import sys
import zipfile

HERE = os.path.dirname(os.path.abspath(__file__))
REF_MODEL = "/root/reference/STEM-GNN/model"
OUT = os.path.join(HERE, "_ref", "stemgnn_reference_model.zip")
FILES = ("__init__.py", "vq.py", "encoder.py", "pt_model.py", "ft_model.py")


def digest(data):
    return hashlib.sha256(data).hexdigest()[:16]


def make(force=False):
    """Returns the archive path, or None when the reference checkout is not on this machine."""
    if not os.path.isdir(REF_MODEL):
        return OUT if os.path.exists(OUT) else None
    srcs = [os.path.join(REF_MODEL, f) for f in FILES]
    if not force and os.path.exists(OUT) and all(os.path.getmtime(s) <= os.path.getmtime(OUT) for s in srcs):
        return OUT
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    manifest = []
    with zipfile.ZipFile(OUT, "w", zipfile.ZIP_DEFLATED) as z:
        for f, s in zip(FILES, srcs):
            data = open(s, "rb").read()
            z.writestr(zipfile.ZipInfo("model/" + f, date_time=(2020, 1, 1, 0, 0, 0)), data)
            manifest.append(f"{digest(data)}  model/{f}  {len(data)} bytes")
        z.writestr(zipfile.ZipInfo("MANIFEST.txt", date_time=(2020, 1, 1, 0, 0, 0)),
                   "unmodified copies of /root/reference/STEM-GNN/model/*.py (sha256[:16], path, size)\n"
                   + "\n".join(manifest) + "\n")
    return OUT


if __name__ == "__main__":
    p = make(force="--force" in sys.argv)
    print(p if p else "oracle/make_ref.py: /root/reference is not on this machine and no archive exists")
