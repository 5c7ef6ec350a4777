from . import adam, lamb
