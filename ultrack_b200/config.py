"""Mirror of the configuration models the segment()/link() path reads.

Same field names, defaults, TOML aliases and validators as the reference
(ultrack/config/segmentationconfig.py:19-107, ultrack/config/config.py:23-88,
ultrack/config/dataconfig.py:25-127).  The reference's own module cannot be imported
without higra (`config/segmentationconfig.py:3`), so `ws_hierarchy` resolves to the real
`hg.watershed_hierarchy_by_*` callable when higra is importable and otherwise to a
named placeholder; either way the path only ever reads its name (`ws_hierarchy_name`).
A real `ultrack.config.MainConfig` instance is accepted everywhere as well (duck typing).
"""
from __future__ import annotations

import logging
import uuid
from enum import Enum
from pathlib import Path
from typing import Any, Callable, Dict, Literal, Optional, Union

import toml
from pydantic import BaseModel, ConfigDict, Field, field_serializer, field_validator, model_validator

LOG = logging.getLogger(__name__)

WS_HIERARCHY_NAMES = ("area", "dynamics", "volume")


class _WsHierarchy:
    """Stand-in for `hg.watershed_hierarchy_by_<name>` when higra is not installed."""

    def __init__(self, name: str):
        self.name = name
        self.__name__ = f"watershed_hierarchy_by_{name}"

    def __call__(self, *args, **kwargs):
        raise RuntimeError(
            f"{self.__name__} is evaluated by the CUDA library (ulk_segment_frame), it is not a "
            "callable CPU routine in ultrack_b200.")

    def __eq__(self, other) -> bool:
        return getattr(other, "__name__", None) == self.__name__

    def __hash__(self) -> int:
        return hash(self.__name__)

    def __repr__(self) -> str:
        return f"<{self.__name__}>"


def _name_to_ws_hier() -> Dict[str, Callable]:
    try:
        import higra as hg  # noqa: F401  (reference behaviour when available)

        return {n: getattr(hg, f"watershed_hierarchy_by_{n}") for n in WS_HIERARCHY_NAMES}
    except Exception:
        return {n: _WsHierarchy(n) for n in WS_HIERARCHY_NAMES}


NAME_TO_WS_HIER = _name_to_ws_hier()


def ws_hierarchy_name(fun: Any) -> str:
    """'area' | 'dynamics' | 'volume' from a config value (string, higra function or placeholder)."""
    if isinstance(fun, str):
        name = fun
    else:
        name = getattr(fun, "__name__", "")
        prefix = "watershed_hierarchy_by_"
        name = name[len(prefix):] if name.startswith(prefix) else name
    if name not in WS_HIERARCHY_NAMES:
        raise ValueError(f"`ws_hierarchy` must match {WS_HIERARCHY_NAMES}. Found {fun}")
    return name


class SegmentationConfig(BaseModel):
    """ultrack/config/segmentationconfig.py:19-107"""

    min_area: int = 100
    min_area_factor: float = 4.0
    max_area: int = 1_000_000
    n_workers: int = 1
    min_frontier: float = 0.0
    threshold: float = 0.5
    max_noise: float = 0.0
    random_seed: Literal["frame", "none", None] = "frame"
    ws_hierarchy: Callable = NAME_TO_WS_HIER["area"]
    anisotropy_penalization: float = 0.0

    model_config = ConfigDict(use_enum_values=True, extra="forbid")

    @field_validator("ws_hierarchy", mode="before")
    def ws_name_to_function(cls, value: str) -> Callable:
        if not isinstance(value, str):
            raise ValueError(f"`ws_hierarchy` must be a string. Found {value} of type {type(value)}.")
        if value not in NAME_TO_WS_HIER:
            raise ValueError(f"`ws_hierarchy` must match {NAME_TO_WS_HIER.keys()}. Found {value}")
        return NAME_TO_WS_HIER[value]

    @field_serializer("ws_hierarchy")
    def serialize_ws_hierarchy(self, value: Callable) -> str:
        return ws_hierarchy_name(value)

    @model_validator(mode="after")
    def area_validator(self) -> "SegmentationConfig":
        if self.min_area > self.max_area:
            raise ValueError("`min_area` must be lower than `max_area`."
                             f"Found min_area={self.min_area}, max_area={self.max_area}")
        return self


class LinkingConfig(BaseModel):
    """ultrack/config/config.py:23-49"""

    max_distance: float = 15.0
    n_workers: int = 1
    max_neighbors: int = 5
    distance_weight: float = 0.0
    z_score_threshold: float = 5.0

    model_config = ConfigDict(extra="forbid")


class TrackingConfig(BaseModel):
    """Pass-through: the tracking (ILP) stage is the unchanged CPU consumer of this path
    (ultrack/config/trackingconfig.py); its fields are kept verbatim for TOML round trips."""

    model_config = ConfigDict(extra="allow")


class DatabaseChoices(Enum):
    sqlite = "sqlite"
    postgresql = "postgresql"
    memory = "memory"


class DataConfig(BaseModel):
    """ultrack/config/dataconfig.py:25-127"""

    n_workers: int = 1
    working_dir: Path = Path(".")
    database_file_name: str = "data.db"
    database: DatabaseChoices = "sqlite"
    address: Optional[str] = None
    in_memory_db_id: int = 0

    model_config = ConfigDict(validate_assignment=True, use_enum_values=True, extra="forbid")

    @field_validator("working_dir")
    def validate_working_dir_writeable(cls, value: Path) -> Path:
        value.mkdir(exist_ok=True, parents=True)
        try:
            tmp_path = value / f".write_test_{uuid.uuid4().hex}"
            open(tmp_path, "w").close()
            tmp_path.unlink()
        except OSError:
            raise ValueError(f"Working directory {value} isn't writable.")
        return value

    @field_serializer("working_dir")
    def serialize_working_dir(self, value: Path) -> str:
        return str(value)

    @model_validator(mode="after")
    def validate_postgresql_parameters(self) -> "DataConfig":
        if self.database == DatabaseChoices.postgresql.value and self.address is None:
            raise ValueError("`data.address` must be defined when `data.database` = `postgresql`."
                             "For example: postgres@localhost:12345/example")
        return self

    @property
    def database_path(self) -> str:
        if self.database == DatabaseChoices.sqlite.value:
            return f"sqlite:///{self.working_dir.absolute()}/{self.database_file_name}"
        elif self.database == DatabaseChoices.memory.value:
            return f"sqlite:///file:{self.in_memory_db_id}?mode=memory&cache=shared&uri=true"
        elif self.database == DatabaseChoices.postgresql.value:
            return f"postgresql://{self.address}"
        raise NotImplementedError(f"Dataset type {self.database} support not implemented.")

    @property
    def metadata_path(self) -> Path:
        return self.working_dir / "metadata.toml"

    def metadata_add(self, data: Dict[str, Any]) -> None:
        metadata = self.metadata
        metadata.update(data)
        LOG.info(f"Updated metadata. New content {metadata}.")
        with open(self.metadata_path, mode="w") as f:
            toml.dump(metadata, f)

    @property
    def metadata(self) -> Dict[str, Any]:
        if not self.metadata_path.exists():
            return {}
        with open(self.metadata_path) as f:
            return toml.load(f)


class MainConfig(BaseModel):
    """ultrack/config/config.py:52-74"""

    data_config: Optional[DataConfig] = Field(default_factory=DataConfig, alias="data")
    segmentation_config: SegmentationConfig = Field(default_factory=SegmentationConfig, alias="segmentation")
    linking_config: LinkingConfig = Field(default_factory=LinkingConfig, alias="linking")
    tracking_config: TrackingConfig = Field(default_factory=TrackingConfig, alias="tracking")


def load_config(path: Union[str, Path]) -> MainConfig:
    """Creates MainConfig from TOML file (ultrack/config/config.py:77-82)."""
    with open(path) as f:
        data = toml.load(f)
        LOG.info(data)
        return MainConfig.model_validate(data)


def save_config(config: MainConfig, path: Union[str, Path]) -> None:
    """Saves MainConfig to TOML file (ultrack/config/config.py:85-88)."""
    with open(path, mode="w") as f:
        toml.dump(config.model_dump(by_alias=True), f)
