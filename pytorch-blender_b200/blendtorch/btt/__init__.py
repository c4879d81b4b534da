"""blendtorch.btt -- B200-native ingest path with the reference's surface.

Same public names as pkg_pytorch/blendtorch/btt/__init__.py for the ingest path
(`RemoteIterableDataset`, `FileDataset`, `FileRecorder`, `FileReader`); the launcher and the gym
environments of the reference are out of scope here and not provided (`DuplexChannel` is).
New, additive names: `DeviceLoader`, `Preprocess`, `preprocess_images`, `FlipRows`, `flip_rows`, `Camera`,
`DataPublisher`, `NativePublisher`, `shard`.
"""
from .constants import DEFAULT_TIMEOUTMS
from .dataset import FileDataset, RemoteIterableDataset, SingleFileDataset
from .file import FileReader, FileRecorder
from . import native
from .preprocess import FlipRows, Preprocess, flip_rows, preprocess_images
from .camera import Camera, Projection
from .loader import DeviceLoader, shard
from .raster import render_cubes
from .publisher import DataPublisher, NativePublisher
from .duplex import DuplexChannel

__version__ = "0.2.0+b200.1"

__all__ = ["RemoteIterableDataset", "FileDataset", "SingleFileDataset", "FileRecorder", "FileReader",
           "DeviceLoader", "Preprocess", "preprocess_images", "FlipRows", "flip_rows", "Camera", "Projection", "DataPublisher", "NativePublisher", "DuplexChannel", "shard", "render_cubes",
           "DEFAULT_TIMEOUTMS", "native"]
