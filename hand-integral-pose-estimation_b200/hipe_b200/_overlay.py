"""Overlay loading for the drop-in modules `augment` and `nrsfm.nrsfm_modules`.

INTEGRATION.md puts this package's directory in FRONT of the reference's on sys.path, so a module
here that has the same name as a reference module is the one Python imports.  The reference's
modules hold far more than the hot path (image / patch augmentation in common/augment.py:106-574,
the sparse-coding layers in procrustes_encoding/nrsfm/nrsfm_modules.py:70-230), and its other code
imports those names from the same module.  So a drop-in module first executes the REFERENCE's own
source of that module, found further down sys.path, inside its own namespace, and then defines the
hot-path functions on top: every other name keeps the reference's behaviour.  Without a reference
tree on the path the drop-in module simply carries the hot-path functions alone."""
import os
import sys


def find_shadowed(own_file, rel_path):
    """Path of the next `rel_path` (e.g. "augment.py", "nrsfm/nrsfm_modules.py") on sys.path that is
    not the file `own_file`; None if there is none."""
    own = os.path.realpath(own_file)
    for d in sys.path:
        if not d:
            d = os.getcwd()
        cand = os.path.join(d, rel_path)
        if os.path.isfile(cand) and os.path.realpath(cand) != own:
            return cand
    return None


def exec_shadowed(module_globals, own_file, rel_path):
    """Execute the shadowed reference module's source in `module_globals`.  Returns
    (path, None) on success, (path, exception) if the reference module could not be executed
    (an import of its own that is unavailable), (None, None) if nothing is shadowed."""
    path = find_shadowed(own_file, rel_path)
    if path is None:
        return None, None
    with open(path, "r") as f:
        code = compile(f.read(), path, "exec")
    keep = {k: module_globals.get(k) for k in ("__file__", "__doc__")}
    try:
        exec(code, module_globals)
    except Exception as e:                       # e.g. matplotlib / torch_batch_svd not installed
        module_globals.update(keep)
        return path, e
    module_globals.update(keep)
    return path, None
