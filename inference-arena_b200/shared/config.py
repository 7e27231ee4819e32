"""Access to experiment.yaml — same interface and error behaviour as the reference's
``src/shared/config.py`` (:57-80 ``get_config``, :83-93 ``reload_config``, :100-138
``get_controlled_variable``) for the keys the processing path reads at import.

Search order: ``$IA_EXPERIMENT_YAML``; the reference's two locations
(``<package>/../../experiment.yaml`` and ``<cwd>/experiment.yaml``, config.py:35-38), so
that inside the reference's tree or containers its own file wins; then the copy shipped
next to this package.
"""

from __future__ import annotations

import os
from functools import lru_cache
from pathlib import Path
from typing import Any

import yaml

_PKG = Path(__file__).resolve().parent


def _candidates() -> list[Path]:
    paths = []
    if os.environ.get("IA_EXPERIMENT_YAML"):
        paths.append(Path(os.environ["IA_EXPERIMENT_YAML"]))
    paths += [_PKG.parent.parent / "experiment.yaml", Path.cwd() / "experiment.yaml",
              _PKG.parent / "experiment.yaml"]
    return paths


def _locate() -> Path:
    cands = _candidates()
    for p in cands:
        if p.exists():
            return p
    return cands[0]


_CONFIG_PATH = _locate()


@lru_cache(maxsize=1)
def get_config() -> dict[str, Any]:
    if not _CONFIG_PATH.exists():
        raise FileNotFoundError(
            f"Experiment configuration not found: {_CONFIG_PATH}\n"
            f"Expected location: {_CONFIG_PATH.absolute()}"
        )
    with open(_CONFIG_PATH) as fh:
        return yaml.safe_load(fh)


def reload_config() -> dict[str, Any]:
    get_config.cache_clear()
    return get_config()


def get_controlled_variable(section: str, key: str) -> Any:
    controlled = get_config().get("controlled_variables", {})
    if section not in controlled:
        raise KeyError(
            f"Section '{section}' not found in controlled_variables. "
            f"Available sections: {list(controlled.keys())}"
        )
    data = controlled[section]
    if key not in data:
        raise KeyError(
            f"Key '{key}' not found in controlled_variables.{section}. "
            f"Available keys: {list(data.keys())}"
        )
    return data[key]


def get_model_config(name: str) -> dict[str, Any]:
    return get_controlled_variable("models", name)
