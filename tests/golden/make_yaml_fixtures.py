"""Snapshot of the reference's own env configs as a JSON fixture (the GPU box has no /root/reference).

    python tests/golden/make_yaml_fixtures.py      (build container only)

Reads the YAML files the reference ships for the Quadrotor env -- competition/level0-3.yaml and getting_started.yaml,
tests/scripts/verbose_api.yaml (quadrotor_config block), tests/scripts/tracking.yaml and the registered default config
safe_control_gym/envs/gym_pybullet_drones/quadrotor.yaml -- and stores their `quadrotor_config` dicts UNCHANGED under
tests/golden/ref_yaml_configs.json.  tests/test_yaml_dropin.py checks the fixture against the files whenever the reference tree
is present and replays the reference scripts' call sequences on it."""
import json
import os

import yaml

REF = "/root/reference"
FILES = {"level0": "competition/level0.yaml", "level1": "competition/level1.yaml", "level2": "competition/level2.yaml",
         "level3": "competition/level3.yaml", "getting_started": "competition/getting_started.yaml",
         "verbose_api": "tests/scripts/verbose_api.yaml", "tracking": "tests/scripts/tracking.yaml",
         "quadrotor_default": "safe_control_gym/envs/gym_pybullet_drones/quadrotor.yaml"}


def read_all(ref=REF):
    out = {}
    for name, rel in FILES.items():
        doc = yaml.safe_load(open(os.path.join(ref, rel)))
        out[name] = {"file": rel, "config": doc.get("quadrotor_config", doc)}
    return out


if __name__ == "__main__":
    here = os.path.dirname(os.path.abspath(__file__))
    with open(os.path.join(here, "ref_yaml_configs.json"), "w") as f:
        json.dump(read_all(), f, indent=1, sort_keys=True)
    print("wrote", len(FILES), "configs")
