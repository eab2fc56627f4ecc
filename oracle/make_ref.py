"""Recipe for ``oracle/_ref``: a verbatim, git-ignored copy of the reference's own Python implementation of the
path (``src/models/*.py``, ``src/loss.py``, ``src/metrics.py``) taken from ``/root/reference`` where that tree
is present (the build container).  TEST / BASELINE INFRASTRUCTURE ONLY.

The reference is a Python repo: there is nothing to compile, the "build" is a copy.  ``oracle/_ref/`` is listed
in ``.gitignore`` (reference sources never enter this repository's history) but not in ``.gpurunignore``, so the
copy travels to the GPU box with the snapshot, where ``/root/reference`` does not exist.  ``bench.py --impl
reference`` and its ``cpu_baseline`` leg time these unmodified modules (``kind: "reference"``) when the copy is
there and fall back to the restatement ``oracle/msgat_oracle.py`` (``kind: "port"``) when it is not.
"""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("MSGAT_REFERENCE", "/root/reference/src")
OUT = os.path.join(HERE, "_ref")
FILES = ["models/__init__.py", "models/attention.py", "models/embeddings.py", "models/msgat.py", "loss.py", "metrics.py"]
EXTRA = ["engine.py"]   # the reference's train / validate loop: drives the drop-in in tests/test_gpu_engine.py


def make(force: bool = False) -> bool:
    """Copies the files; returns True when oracle/_ref is usable afterwards."""
    if not os.path.isdir(REF):
        return available()
    for f in FILES + EXTRA:
        src, dst = os.path.join(REF, f), os.path.join(OUT, f)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        if force or not os.path.exists(dst) or os.path.getmtime(dst) < os.path.getmtime(src):
            shutil.copyfile(src, dst)
    return available()


def available() -> bool:
    return all(os.path.exists(os.path.join(OUT, f)) for f in FILES)


def load():
    """-> (models package, loss module, metrics module) of the unmodified reference, imported from oracle/_ref."""
    if not available():
        raise RuntimeError("oracle/_ref is missing: run oracle/make_ref.py where /root/reference exists")
    sys.dont_write_bytecode = True
    if OUT not in sys.path:
        sys.path.insert(0, OUT)
    import importlib
    models = importlib.import_module("models")
    loss = importlib.import_module("loss")
    metrics = importlib.import_module("metrics")
    for mod in (models, loss, metrics):
        if not os.path.abspath(mod.__file__).startswith(OUT):
            raise RuntimeError(f"{mod.__name__} was imported from {mod.__file__}, not from oracle/_ref")
    return models, loss, metrics


def load_engine():
    """-> the unmodified reference ``engine`` module (imports ``loss`` / ``metrics`` from oracle/_ref), or None."""
    if not available() or not os.path.exists(os.path.join(OUT, "engine.py")):
        return None
    load()
    import importlib
    return importlib.import_module("engine")


if __name__ == "__main__":
    print("oracle/_ref available:", make(force="-f" in sys.argv))
