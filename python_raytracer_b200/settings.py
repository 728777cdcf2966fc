"""The `data.settings` subset read by the trace path, with the reference's parsing rules.

Reference: data.py:15-68.  Every key is mandatory in config.cfg (cfg.get* raises when missing) and
``value or default`` replaces *falsy* values only (so ``samples = 0`` becomes 1, ``fov = 0`` becomes 90).
A live reference ``data.settings`` object can be wrapped directly with ``Settings.from_object`` -- the host
integration reuses the engine's own parsed settings instead of re-reading the file.
"""
from __future__ import annotations

import configparser
import os

from .abi import VtSettings

# name -> (section, getter, default), restricted to what Camera.trace / Camera.tile / chunk_update read
_FIELDS = {
    "width": ("WINDOW", "getint", 64), "height": ("WINDOW", "getint", 64),
    "culling": ("RENDER", "getboolean", False), "static": ("RENDER", "getboolean", False),
    "samples": ("RENDER", "getint", 1), "shutter": ("RENDER", "getfloat", 0), "fov": ("RENDER", "getfloat", 90),
    "falloff": ("RENDER", "getfloat", 0), "chunk_size": ("RENDER", "getint", 16), "chunk_lod": ("RENDER", "getint", 0),
    "dof": ("RENDER", "getfloat", 0), "dist_min": ("RENDER", "getint", 0), "dist_max": ("RENDER", "getint", 32),
    "max_light": ("RENDER", "getfloat", 0), "max_bounces": ("RENDER", "getfloat", 0),
    "lod_bounces": ("RENDER", "getfloat", 0), "lod_samples": ("RENDER", "getfloat", 0),
    "lod_random": ("RENDER", "getfloat", 0), "lod_edge": ("RENDER", "getfloat", 0), "threads": ("RENDER", "getint", 1),
}


_SORTED_FIELDS = tuple(sorted(_FIELDS))


class Settings:
    """Attribute bag with the reference's names.  Derived fields follow data.py:64-68."""

    def __init__(self, **kw):
        for name, (_, _, default) in _FIELDS.items():
            setattr(self, name, default)
        for k, v in kw.items():
            setattr(self, k, v)

    @classmethod
    def from_config(cls, path: str) -> "Settings":
        """Parse mods/<mod>/config.cfg like data.py:16-63 (missing keys raise, falsy values take the default)."""
        if not os.path.isfile(path):
            raise FileNotFoundError(path)
        cfg = configparser.RawConfigParser()
        cfg.read(path)
        s = cls()
        for name, (section, getter, default) in _FIELDS.items():
            setattr(s, name, getattr(cfg, getter)(section, name) or default)
        return s

    @classmethod
    def from_object(cls, obj) -> "Settings":
        """Wrap a reference `data.settings` store or a plain dict."""
        get = (lambda k, d: obj.get(k, d)) if isinstance(obj, dict) else (lambda k, d: getattr(obj, k, d))
        s = cls()
        for name, (_, _, default) in _FIELDS.items():
            setattr(s, name, get(name, default))
        return s

    # ---- derived (data.py:64-68) ----
    @property
    def window(self):
        return (self.width, self.height)

    @property
    def proportions(self):
        return ((self.width + self.height) / 2) / max(self.width, self.height)

    @property
    def chunk_radius(self):
        return round(self.chunk_size / 2)

    def to_struct(self, background: bool = True) -> VtSettings:
        v = VtSettings()
        v.width, v.height, v.samples = int(self.width), int(self.height), int(self.samples)
        v.static_noise = 1 if self.static else 0
        v.chunk_size, v.chunk_radius = int(self.chunk_size), int(self.chunk_radius)
        v.background = 1 if background else 0
        v.reserved = 0
        for n in ("shutter", "falloff", "dof", "dist_min", "dist_max", "max_light", "max_bounces",
                  "lod_bounces", "lod_samples", "lod_random", "lod_edge"):
            setattr(v, n, float(getattr(self, n)))
        v.proportions = float(self.proportions)
        return v

    def key(self):
        return tuple(getattr(self, n) for n in _SORTED_FIELDS)
