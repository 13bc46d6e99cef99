"""Import shim: the package directory is `diffusion-piano_b200/` (a hyphen is not a legal
Python identifier), so `import diffusion_piano_b200` resolves to it through __path__."""
import os as _os

__path__ = [_os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "diffusion-piano_b200")]
with open(_os.path.join(__path__[0], "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(__path__[0], "__init__.py"), "exec"))
