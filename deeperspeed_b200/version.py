__version__ = "0.3.15"
__version_major__, __version_minor__, __version_patch__ = 0, 3, 15
git_hash = "b200-native"
git_branch = "main"
