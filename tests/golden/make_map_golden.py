"""Writes tests/golden/map_server_golden.npz from the reference's own unit-test
material (run in the build container, where /root/reference is mounted):

  * the decoded pixels of map_server/test/{testmap.png, testmap.bmp, testmap2.png,
    spectrum.png} (decoded with PIL; SDL_image hands the loader the same bytes:
    24-bit RGB for the first three, 8-bit grey for the spectrum);
  * the expected OccupancyGrid contents from map_server/test/test_constants.cpp
    (g_valid_image_content, tmap2::g_valid_image_content) and the loader
    parameters the tests use (utest.cpp:48,68; testmap2.yaml);
  * the value counts utest.cpp:121-152 expects for spectrum.png;
  * the pixels of costmap_2d/test/TenByTen.pgm, the static map of costmap_2d's inflation tests.
"""
import os
import re

import numpy as np
from PIL import Image

REF = "/root/reference/map_server/test"
HERE = os.path.dirname(os.path.abspath(__file__))


def content(name, text):
    m = re.search(r"const char " + name + r"\[\] = \{(.*?)\};", text, re.S)
    return np.array([int(v) for v in m.group(1).replace("\n", " ").split(",") if v.strip()], dtype=np.int8)


def main():
    text = open(os.path.join(REF, "test_constants.cpp")).read()
    first, second = text.split("namespace tmap2")
    out = {}
    for f in ("testmap.png", "testmap.bmp", "testmap2.png", "spectrum.png"):
        out["pixels_" + f.replace(".", "_")] = np.array(Image.open(os.path.join(REF, f)))
    # costmap_2d's static test map (costmap_2d/test/TenByTen.pgm + .yaml: negate 0, occupied 0.65, free 0.196),
    # the map under inflation_tests.cpp's testInflation / testInflation2
    out["pixels_tenbyten_pgm"] = np.array(Image.open("/root/reference/costmap_2d/test/TenByTen.pgm"))
    out["params_tenbyten"] = np.array([0, 0.65, 0.196])
    out["expect_testmap"] = content("g_valid_image_content", first).reshape(10, 10)
    out["expect_testmap2"] = content("g_valid_image_content", second).reshape(12, 12)
    # utest.cpp:48,68 : negate false, occ 0.65, free 0.1, default mode TRINARY
    out["params_testmap"] = np.array([0, 0.65, 0.1])
    # testmap2.yaml : negate 1, occupied_thresh 0.5, free_thresh 0.2
    out["params_testmap2"] = np.array([1, 0.5, 0.2])
    # utest.cpp:126-151 : spectrum.png with occ 0.65, free 0.1
    out["spectrum_trinary_counts"] = np.array([90, 26, 140])   # value 100, 0, 255(-1)
    np.savez_compressed(os.path.join(HERE, "map_server_golden.npz"), **out)
    print({k: getattr(v, "shape", v) for k, v in out.items()})


if __name__ == "__main__":
    main()
