import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def port_oracle():
    from oracle import oracle as O

    return O.port()


@pytest.fixture(scope="session")
def ref_oracle():
    from oracle import oracle as O

    r = O.ref()
    if r is None:
        pytest.skip("oracle/_ref not built here (needs /root/reference; `make -C oracle ref`)")
    return r


@pytest.fixture(scope="session")
def best_oracle():
    from oracle import oracle as O

    return O.best()


@pytest.fixture(scope="session")
def scene_cache():
    from caveengine_b200 import scenes

    cache = {}

    def get(name):
        if name not in cache:
            cache[name] = {"tiny": scenes.tiny_scene, "cornell": scenes.cornell_scene, "heightfield": scenes.heightfield_scene, "sponza": scenes.sponza_scene,
                           "breakfast": scenes.breakfast_scene,  # BASELINE config 4's scene (1,048,576 triangles)
                           "animated": lambda: scenes.animated_scene(500, 500),  # config 5's scene at test size (500,000 + 128 triangles)
                           }[name]()
        return cache[name]

    return get
