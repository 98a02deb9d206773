"""heat3d-b200: B200-native heat3D hot path (C-ABI library + C++ driver) and its Python test mirror.

The directory name carries hyphens (it mirrors the reference repository's name), so import it through
`__graft_entry__.load_package()` or put this directory on sys.path and `import heat3d`.
"""
from .heat3d import *  # noqa: F401,F403
