"""TEST INFRASTRUCTURE -- import the UNMODIFIED reference from /root/reference, or, on a box
without it (the GPU box), from the staged copy oracle/build_ref.sh left in oracle/_ref/.

It is used by tests/golden/make_golden.py to generate the committed golden vectors, by the CPU
tests that pin oracle/ against the reference when the reference tree is present, and by
bench.py's reference arms (the reference's own CPU path, and the unmodified reference on the
same GPU) -- never by the product.

No reference source is copied: the reference's Python is imported in place, and the two
native modules it lazily imports with a relative import
(reference: DCVC-DC/src/models/entropy_models.py:13,19) are satisfied by pre-registering
the objects built by oracle/build_ref.sh in sys.modules under the names the relative import
resolves to.
"""
import importlib
import importlib.util
import os
import sys
import sysconfig
import types

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.environ.get("DCVC_REFERENCE_ROOT", "/root/reference")
_SUF = sysconfig.get_config_var("EXT_SUFFIX")
# where /root/reference does not exist (the GPU box) the staged copy made by build_ref.sh in the
# build container (oracle/_ref/<project>/, git-ignored, travels with gpurun) is imported instead
STAGED_ROOT = os.path.join(HERE, "_ref")


def project_root(project="DCVC-DC"):
    for root in (REF_ROOT, STAGED_ROOT):
        if os.path.isdir(os.path.join(root, project, "src", "models")):
            return os.path.join(root, project)
    return None


def reference_available(project="DCVC-DC"):
    return project_root(project) is not None and \
        os.path.isfile(os.path.join(HERE, "_ref", "MLCodec_rans" + _SUF))


def _load_ext(modname, path, register_as):
    spec = importlib.util.spec_from_file_location(modname, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    for name in register_as:
        sys.modules[name] = mod
    return mod


def load_native(project="DCVC-DC"):
    """Load oracle/_ref/*.so (the reference's own C++ compiled here) as python modules."""
    if project == "DCVC-DC":
        rans = _load_ext("MLCodec_rans", os.path.join(HERE, "_ref", "MLCodec_rans" + _SUF),
                         ["src.models.MLCodec_rans"])
        cxx = _load_ext("MLCodec_CXX", os.path.join(HERE, "_ref", "MLCodec_CXX" + _SUF),
                        ["src.models.MLCodec_CXX"])
    else:
        rans = _load_ext("MLCodec_rans", os.path.join(HERE, "_ref", "hem", "MLCodec_rans" + _SUF),
                         ["src.entropy_models.MLCodec_rans"])
        cxx = _load_ext("MLCodec_CXX", os.path.join(HERE, "_ref", "MLCodec_CXX" + _SUF),
                        ["src.entropy_models.MLCodec_CXX"])
    return rans, cxx


def _install_shims():
    # pytorch_msssim is imported unconditionally by the reference harness
    # (DCVC-DC/test_video.py:23, DCVC-HEM/src/models/common_model.py:9) but absent here.
    if "pytorch_msssim" not in sys.modules:
        try:
            importlib.import_module("pytorch_msssim")
        except ImportError:
            stub = types.ModuleType("pytorch_msssim")

            def ms_ssim(*a, **k):
                raise RuntimeError("pytorch_msssim stub: MS-SSIM not available")

            class MS_SSIM:  # noqa: N801
                def __init__(self, *a, **k):
                    pass

                def __call__(self, x, *a, **k):
                    # DCVC-HEM's forward() evaluates self.ssim unconditionally
                    # (image_model.py:80, video_model.py:454): a zero stands in, it enters
                    # no bitstream, bit count or reconstruction
                    import torch
                    return torch.zeros(x.shape[0])
            stub.ms_ssim = ms_ssim
            stub.MS_SSIM = MS_SSIM
            sys.modules["pytorch_msssim"] = stub
    # DMC.encode_decode calls torch.cuda.synchronize unconditionally
    # (DCVC-DC/src/models/video_model.py:528,533,539).
    import torch
    if not torch.cuda.is_available():
        torch.cuda.synchronize = lambda *a, **k: None


def import_reference(project="DCVC-DC"):
    """Returns the reference `src` package namespace for one project (DC or HEM).

    DC and HEM both call their top-level package `src`; only one may live in a process.
    """
    if not reference_available(project):
        raise RuntimeError("reference tree or oracle/_ref not available")
    root = project_root(project)
    if "src" in sys.modules:
        cur = getattr(sys.modules["src"], "__path__", [""])
        if not any(os.path.abspath(p).startswith(root) for p in cur):
            raise RuntimeError("another reference project is already imported as `src`")
    _install_shims()
    if root not in sys.path:
        sys.path.insert(0, root)
    importlib.import_module("src")
    load_native(project)
    ns = types.SimpleNamespace()
    if project == "DCVC-DC":
        ns.video_model = importlib.import_module("src.models.video_model")
        ns.image_model = importlib.import_module("src.models.image_model")
        ns.entropy_models = importlib.import_module("src.models.entropy_models")
        ns.stream_helper = importlib.import_module("src.utils.stream_helper")
        ns.DMC = ns.video_model.DMC
        ns.IntraNoAR = ns.image_model.IntraNoAR
    else:
        ns.video_model = importlib.import_module("src.models.video_model")
        ns.image_model = importlib.import_module("src.models.image_model")
        ns.entropy_models = importlib.import_module("src.entropy_models.entropy_models")
        ns.DMC = ns.video_model.DMC
        ns.IntraNoAR = ns.image_model.IntraNoAR
    return ns
