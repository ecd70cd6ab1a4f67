"""Import-time stand-in for the `sgl_kernel` wheel (absent from this image; test scaffolding only).

The reference imports `sgl_kernel` names at module scope (unconditionally in a few files the unit-test kit pulls
in, and under `if is_cuda():` in many more) although the dense-attention cases never call them.  Every attribute
and sub-module resolves to a placeholder object: attribute access yields another placeholder, *calling* one
raises -- an accidental use fails loudly instead of computing anything."""
import importlib.abc
import importlib.machinery
import sys
import types


class _Missing:
    def __init__(self, qualname):
        object.__setattr__(self, "_qualname", qualname)

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _Missing(f"{self._qualname}.{name}")

    def __call__(self, *args, **kwargs):
        raise NotImplementedError(f"{self._qualname}: sgl_kernel is not installed in this image (stub)")

    def __repr__(self):
        return f"<missing {self._qualname}>"


class _StubModule(types.ModuleType):
    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _Missing(f"{self.__name__}.{name}")


class _Finder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    def find_spec(self, fullname, path=None, target=None):
        if fullname.startswith("sgl_kernel."):
            return importlib.machinery.ModuleSpec(fullname, self, is_package=True)
        return None

    def create_module(self, spec):
        m = _StubModule(spec.name)
        m.__path__ = []
        return m

    def exec_module(self, module):
        if module.__name__ == "sgl_kernel.utils":
            # a capability probe, not a kernel: the reference's Triton backend asks it at construction
            # (triton_backend.py:231); without the wheel there is no PDL launch path
            module.is_arch_support_pdl = lambda: False


sys.meta_path.append(_Finder())
__version__ = "0.0.0+stub"


def fast_topk(values, topk, dim):
    """The one real function: the wheel's `fast_topk` is plain torch (max for k = 1, else topk); the reference's
    frozen-KV MTP draft worker calls it between attention steps in one of the kit's runner tests."""
    import torch

    if topk == 1:
        return torch.max(values, dim=dim, keepdim=True)
    return torch.topk(values, topk, dim=dim)


def __getattr__(name):
    if name.startswith("__"):
        raise AttributeError(name)
    return _Missing(f"sgl_kernel.{name}")
