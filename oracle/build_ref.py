"""Recipe: stage the reference's OWN Python modules of the hot path under ``oracle/_ref/`` (git-ignored).

    python -m oracle.build_ref            # in the build container, where /root/reference exists

TEST / BASELINE INFRASTRUCTURE (see oracle/__init__.py).  The reference is pure Python: there is
nothing to compile, so "building" ``oracle/_ref`` is a verbatim copy of the files the path runs through, taken from
where they lie under ``/root/reference`` into a directory that is kept out of git history (``.gitignore``) but
travels to the GPU box with the snapshot, like the built ``.so``.  Nothing under ``vae_npvc_b200/`` imports it.
With it present,

  * ``bench.py --impl reference`` and the ``cpu_baseline`` leg time the reference's own ``vae_npvc.model.vqvae.Model``
    with ``trainer/basic.py:55-79`` around it (``kind: "reference"``) instead of the oracle port;
  * ``bench.py --impl reference --ref-device cuda`` runs the same stock modules on the B200 through eager PyTorch
    (cuDNN / cuBLAS) - the "existing kernel on the same box" of BASELINE.md section 4;
  * GPU tests can compare with the reference classes directly.

Files: vae_npvc/model/{vqvae,layers,layers_vq,vqvae2,layers_gst}.py, vae_npvc/trainer/{basic,radam}.py.
A manifest with the sha256 of every staged file is written beside them.
"""
import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.environ.get('VNB_REFERENCE_ROOT', '/root/reference')
OUT = os.path.join(HERE, '_ref')
FILES = [
    'vae_npvc/model/vqvae.py', 'vae_npvc/model/layers.py', 'vae_npvc/model/layers_vq.py',
    'vae_npvc/model/vqvae2.py', 'vae_npvc/model/layers_gst.py',
    'vae_npvc/trainer/basic.py', 'vae_npvc/trainer/radam.py',
]


def available():
    """True when the staged reference modules can be imported (oracle/_ref present)."""
    return os.path.exists(os.path.join(OUT, 'vae_npvc', 'model', 'vqvae.py'))


def build(force=False):
    """Copy the reference's modules into oracle/_ref.  No-op (returns False) when /root/reference is absent -
    the GPU box only uses what was staged in the build container."""
    if not os.path.isdir(REF_ROOT):
        return False
    manifest = {}
    for rel in FILES:
        src, dst = os.path.join(REF_ROOT, rel), os.path.join(OUT, rel)
        if not os.path.exists(src):
            continue
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        if force or not os.path.exists(dst) or open(src, 'rb').read() != open(dst, 'rb').read():
            shutil.copyfile(src, dst)
        manifest[rel] = hashlib.sha256(open(dst, 'rb').read()).hexdigest()
    with open(os.path.join(OUT, 'MANIFEST.json'), 'w') as f:
        json.dump({'source': REF_ROOT, 'files': manifest}, f, indent=1, sort_keys=True)
    return True


def import_reference():
    """-> the reference's ``vae_npvc`` namespace package (from oracle/_ref, else from /root/reference).
    Raises ImportError when neither exists."""
    root = OUT if available() else (REF_ROOT if os.path.isdir(os.path.join(REF_ROOT, 'vae_npvc')) else None)
    if root is None:
        raise ImportError('the reference modules are not staged (run `python -m oracle.build_ref` in the build '
                          'container) and /root/reference is absent')
    if root not in sys.path:
        sys.path.insert(0, root)
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        import vae_npvc.model.vqvae  # noqa: F401  (namespace package: no __init__.py)
    import vae_npvc
    return vae_npvc


if __name__ == '__main__':
    ok = build(force='--force' in sys.argv)
    print('staged %d reference files under %s' % (len(FILES), OUT) if ok else 'no reference tree at %s' % REF_ROOT)
