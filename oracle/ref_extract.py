"""TEST INFRASTRUCTURE ONLY — build step of oracle/_ref: copies selected FUNCTION DEFINITIONS, verbatim, out of the
reference's camera-model sources (where they lie under /root/reference) into generated translation units under
oracle/_ref/gen/ (git-ignored), one per reference file.  No reference source is stored in this repository.

Why not compile the .cc files whole: besides ``liftProjective`` they hold calibration / undistortion-map /
YAML code that needs the real OpenCV and Eigen SDKs (cv::FileStorage, cv::solvePnP, Eigen::JacobiSVD, …), absent
here.  The functions on the tracker's path are plain double arithmetic; they are taken by NAME (signature regex +
brace matching, so the recipe survives line shifts) and compiled against the reference's own headers plus the
stand-ins in oracle/ref_shim/.  The functions picked are listed in WANT below; every other virtual gets an aborting
stub in oracle/ref_driver.cpp.

usage: python oracle/ref_extract.py /root/reference oracle/_ref/gen
"""
from __future__ import annotations

import os
import re
import sys

CAMERA_DIR = "src/common/camera_models"

# file -> regexes matched against the (whitespace-normalised) signature of a top-level definition
WANT = {
    "Camera.cc": [
        r"^Camera::Parameters::Parameters\(",
        r"Camera::Parameters::(modelType|cameraName|imageWidth|imageHeight|nIntrinsics)\(void\)",
        r"Camera::mask\(void\)",
    ],
    "CameraFactory.cc": [
        r"^bool iequals\(",
        r"^CameraFactory::CameraFactory\(",
        r"CameraFactory::instance\(void\)",
        r"CameraFactory::generateCameraFromYamlFile\(",
    ],
    "PinholeCamera.cc": [
        r"PinholeCamera::Parameters::readFromYamlFile\(",
        r"^PinholeCamera::Parameters::Parameters\(",
        r"PinholeCamera::Parameters::(k1|k2|p1|p2|fx|fy|cx|cy)\(void\)",
        r"PinholeCamera::Parameters::operator=\(",
        r"^PinholeCamera::PinholeCamera\(",
        r"PinholeCamera::(modelType|cameraName|imageWidth|imageHeight|parameterCount|getParameters)\(void\)",
        r"PinholeCamera::(liftSphere|liftProjective)\(",
        r"PinholeCamera::distortion\(const Eigen::Vector2d& p_u, Eigen::Vector2d& d_u\) const",
        r"PinholeCamera::(setParameters|readParameters|writeParameters)\(",
    ],
    "CataCamera.cc": [
        r"CataCamera::Parameters::readFromYamlFile\(",
        r"^CataCamera::Parameters::Parameters\(",
        r"CataCamera::Parameters::(xi|k1|k2|p1|p2|gamma1|gamma2|u0|v0)\(void\)",
        r"CataCamera::Parameters::operator=\(",
        r"^CataCamera::CataCamera\(",
        r"CataCamera::(modelType|cameraName|imageWidth|imageHeight|parameterCount|getParameters)\(void\)",
        r"CataCamera::(liftSphere|liftProjective)\(",
        r"CataCamera::distortion\(const Eigen::Vector2d& p_u, Eigen::Vector2d& d_u\) const",
        r"CataCamera::(setParameters|readParameters|writeParameters)\(",
    ],
    "EquidistantCamera.cc": [
        r"EquidistantCamera::Parameters::readFromYamlFile\(",
        r"^EquidistantCamera::Parameters::Parameters\(",
        r"EquidistantCamera::Parameters::(k2|k3|k4|k5|mu|mv|u0|v0)\(void\)",
        r"EquidistantCamera::Parameters::operator=\(",
        r"^EquidistantCamera::EquidistantCamera\(",
        r"EquidistantCamera::(modelType|cameraName|imageWidth|imageHeight|parameterCount|getParameters)\(void\)",
        r"EquidistantCamera::(liftSphere|liftProjective|backprojectSymmetric)\(",
        r"EquidistantCamera::(setParameters|readParameters|writeParameters)\(",
    ],
    "ScaramuzzaCamera.cc": [
        r"^bool iequals\(",
        r"OCAMCamera::Parameters::readFromYamlFile\(",
        r"^OCAMCamera::Parameters::Parameters\(",
        r"OCAMCamera::Parameters::operator=\(",
        r"^OCAMCamera::OCAMCamera\(",
        r"OCAMCamera::(modelType|cameraName|imageWidth|imageHeight|parameterCount|getParameters)\(void\)",
        r"OCAMCamera::(liftSphere|liftProjective)\(",
        r"OCAMCamera::(setParameters|readParameters|writeParameters)\(",
    ],
}

_SKIP_START = ("#", "/", "*", "}", "namespace", "using", "template", "typedef", "static", "extern")


def definitions(text: str):
    """Yield (first_line_no, signature, body_text) for every top-level definition (column-0 start, closed by a
    column-0 '}' or on its own line)."""
    lines = text.split("\n")
    i, n = 0, len(lines)
    while i < n:
        ln = lines[i]
        if not ln or ln[0].isspace() or ln.startswith(_SKIP_START):
            i += 1
            continue
        start = i
        # signature may continue on indented lines until the line that opens the body
        sig = []
        while i < n:
            sig.append(lines[i].strip())
            if "{" in lines[i] or lines[i].rstrip().endswith(";"):
                break
            i += 1
        if i >= n or lines[i].rstrip().endswith(";") and "{" not in lines[i]:
            i += 1
            continue
        # body: to the matching column-0 '}', unless the opening line already closes it ('{}' one-liner)
        depth = 0
        j = start
        done = False
        while j < n:
            depth += lines[j].count("{") - lines[j].count("}")
            if j >= i and depth == 0:
                done = True
                break
            j += 1
        if not done:
            break
        yield start + 1, " ".join(sig), "\n".join(lines[start:j + 1])
        i = j + 1


HEADER = ["// GENERATED by oracle/ref_extract.py from the reference's sources — do not commit (oracle/_ref/ is git-ignored).",
          "#include <algorithm>", "#include <cctype>", "#include <cmath>", "#include <cstring>", "#include <iostream>", "#include <string>",
          "#include <vector>",
          '#include "common/camera_models/Camera.h"', '#include "common/camera_models/PinholeCamera.h"',
          '#include "common/camera_models/CataCamera.h"', '#include "common/camera_models/EquidistantCamera.h"',
          '#include "common/camera_models/ScaramuzzaCamera.h"', '#include "common/camera_models/CameraFactory.h"',
          "namespace common {", "namespace camera_models {", ""]


def extract(ref_root: str):
    """{reference file name: generated translation unit}, manifest lines.  One TU per reference file, as in the reference
    (file-local helpers such as iequals() exist in more than one of them)."""
    units, manifest = {}, []
    for fname, pats in WANT.items():
        path = os.path.join(ref_root, CAMERA_DIR, fname)
        text = open(path, encoding="utf-8").read()
        regs = [re.compile(p) for p in pats]
        hit = {p: 0 for p in pats}
        out = list(HEADER)
        for line_no, sig, body in definitions(text):
            # "^" patterns match at the start of the signature (constructors and free helpers have no qualified return type)
            for p, r in zip(pats, regs):
                if (r.match(sig) if p.startswith("^") else r.search(sig)):
                    out.append(f"// ---- {CAMERA_DIR}/{fname}:{line_no}")
                    # a file-local helper (anonymous namespace in the reference) stays file-local here
                    out.append("namespace {\n" + body + "\n}" if "::" not in sig.split("(")[0] else body)
                    out.append("")
                    manifest.append(f"{fname}:{line_no}-{line_no + body.count(chr(10))}  {sig.split('{')[0].strip()}")
                    hit[p] += 1
                    break
        missing = [p for p, c in hit.items() if c == 0]
        if missing:
            raise SystemExit(f"ref_extract: nothing in {fname} matches {missing}")
        out += ["}  // namespace camera_models", "}  // namespace common", ""]
        units[fname] = "\n".join(out)
    return units, manifest


def main():
    ref_root, dst_dir = sys.argv[1], sys.argv[2]
    units, manifest = extract(ref_root)
    os.makedirs(dst_dir, exist_ok=True)
    for f in os.listdir(dst_dir):
        if f.startswith("ref_") and f.endswith(".cc"):
            os.unlink(os.path.join(dst_dir, f))
    for fname, src in units.items():
        with open(os.path.join(dst_dir, "ref_" + fname), "w", encoding="utf-8") as f:
            f.write(src)
    with open(os.path.join(dst_dir, "manifest.txt"), "w", encoding="utf-8") as f:
        f.write("\n".join(manifest) + "\n")
    print(f"ref_extract: {len(manifest)} definitions -> {dst_dir}/ref_*.cc")


if __name__ == "__main__":
    main()
