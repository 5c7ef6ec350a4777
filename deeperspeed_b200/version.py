"""Version of the API surface this package provides (DeeperSpeed 0.3.15) and the checkout it was built from."""
import os
import re

__version__ = "0.3.15"
__version_major__, __version_minor__, __version_patch__ = (int(x) for x in re.match(r"(\d+)\.(\d+)\.(\d+)", __version__).groups())


def _checkout_state():
    """(short commit hash, branch) read straight from .git next to the package - no subprocess at import time;
    a tree without history (a snapshot shipped to a GPU box, an sdist) reports itself as such."""
    git_dir = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), ".git")
    try:
        head = open(os.path.join(git_dir, "HEAD")).read().strip()
        if not head.startswith("ref:"):
            return head[:7], "[detached]"
        ref = head.split(None, 1)[1]
        branch = ref.rsplit("/", 1)[-1]
        ref_file = os.path.join(git_dir, ref)
        if os.path.isfile(ref_file):
            return open(ref_file).read().strip()[:7], branch
        for line in open(os.path.join(git_dir, "packed-refs")):
            if line.strip().endswith(" " + ref):
                return line.split()[0][:7], branch
        return "[unborn]", branch
    except OSError:
        return "b200-native", "[snapshot]"


git_hash, git_branch = _checkout_state()
