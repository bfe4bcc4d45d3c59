"""Importable alias of the `image-compression-cpp_b200/` package directory (a hyphenated
directory cannot be imported by name): extends the package path so that
`import image_compression_cpp_b200.reward` resolves to `image-compression-cpp_b200/reward.py`."""
import os as _os

__path__.append(_os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                              "image-compression-cpp_b200"))
