"""Import hook that makes the reference's own import lines resolve to cardioflow_b200 -- without editing the reference.

The reference's ``nnunet``, ``nnunet.lib`` and ``nnunet.network_architecture`` are regular packages (they have
``__init__.py``), so a same-named directory earlier on ``PYTHONPATH`` is never consulted for their sub-modules.  This
finder sits at the front of ``sys.meta_path`` instead and answers for exactly three module names:

    nnunet.lib.raft_initial                    (imported at nnunet/network_architecture/SegFlowGaussian.py:42)
    nnunet.network_architecture.integration    (imported by 10+ files, e.g. SegFlowGaussian.py:35)
    nnunet.lib.raft                            (CorrVolume, SegFlowGaussian.py:41; the file is missing upstream)

with the files under ``shim/nnunet/...``, which re-export the cardioflow_b200 classes and pass every OTHER name
(``BasicUpdateBlock``, ...) through from the module the reference tree itself provides under that name.

Use:  ``PYTHONPATH=/path/to/repo/shim:...`` (``shim/sitecustomize.py`` calls ``install()`` at interpreter start), or
``import cardioflow_dropin; cardioflow_dropin.install()`` before the first ``import nnunet...``.
Pure stdlib: nothing heavy is imported until one of the two modules is.
"""
import importlib.abc
import importlib.util
import os
import sys

SHIM_DIR = os.path.dirname(os.path.abspath(__file__))
REPO_ROOT = os.path.dirname(SHIM_DIR)
TARGETS = {
    "nnunet.lib.raft_initial": os.path.join(SHIM_DIR, "nnunet", "lib", "raft_initial.py"),
    "nnunet.network_architecture.integration": os.path.join(SHIM_DIR, "nnunet", "network_architecture", "integration.py"),
    "nnunet.lib.raft": os.path.join(SHIM_DIR, "nnunet", "lib", "raft.py"),     # CorrVolume (SegFlowGaussian.py:41; parity unpinned)
}


class _Finder(importlib.abc.MetaPathFinder):
    def find_spec(self, fullname, path=None, target=None):
        file = TARGETS.get(fullname)
        if file is None:
            return None
        return importlib.util.spec_from_file_location(fullname, file)


def upstream_module(fullname):
    """The module the reference tree itself would have loaded under ``fullname`` (found on the parent package's
    ``__path__``), executed under a private name; None when the tree has no such file (nnunet/lib/raft_initial.py was
    never committed upstream)."""
    private = "_cardioflow_upstream." + fullname
    if private in sys.modules:
        return sys.modules[private]
    parent_name, _, leaf = fullname.rpartition(".")
    parent = sys.modules.get(parent_name)
    for d in list(getattr(parent, "__path__", []) or []):
        cand = os.path.join(d, leaf + ".py")
        if os.path.isfile(cand) and os.path.abspath(cand) != os.path.abspath(TARGETS[fullname]):
            spec = importlib.util.spec_from_file_location(private, cand)
            mod = importlib.util.module_from_spec(spec)
            mod.__package__ = parent_name          # relative imports inside the upstream file keep working
            sys.modules[private] = mod
            try:
                spec.loader.exec_module(mod)
            except BaseException:
                del sys.modules[private]
                raise
            return mod
    return None


def install():
    if REPO_ROOT not in sys.path:
        sys.path.append(REPO_ROOT)                 # so that `import cardioflow_b200` resolves
    if not any(isinstance(f, _Finder) for f in sys.meta_path):
        sys.meta_path.insert(0, _Finder())


def uninstall():
    sys.meta_path[:] = [f for f in sys.meta_path if not isinstance(f, _Finder)]
