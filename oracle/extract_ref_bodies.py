#!/usr/bin/env python3
"""TEST INFRASTRUCTURE ONLY (build step of `make -C oracle refmatch`).

Copies, at BUILD time, the bodies of the Frame / KeyFrame / MapPoint member functions that lie on the hot path out of the
reference tree into git-ignored oracle/_ref/gen/*.inc, so that they can be compiled — unmodified, as the reference wrote
them — against the shim classes of oracle/ref_shim/match_prelude.h.  The files they live in (src/Frame.cc, src/KeyFrame.cc,
src/MapPoint.cc) cannot be compiled whole here: the rest of those files needs g2o, DBoW2's vocabulary, the IMU code and the
Map.  Nothing is copied into the repository: oracle/_ref/ is in .gitignore.

Each function is found by the text of its definition line and cut at the matching closing brace; a definition that is not
found aborts the build (the reference moved: fix the table, do not guess).
"""
import os
import re
import sys

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
OUT = sys.argv[2] if len(sys.argv) > 2 else os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "gen")

# output file -> list of (source file, regex matching the first line of the definition)
TABLE = {
    # members defined inline in the class bodies (included inside the shim classes)
    "frame_members.inc": [
        ("include/Frame.h", r"inline Vector3f World2Camera\(const Vector3f &p_w, const SE3f &T_c_w\) const \{"),
        ("include/Frame.h", r"inline Vector2f Camera2Pixel\(const Vector3f &p_c\) const \{"),
        ("include/Frame.h", r"Vector2f World2Pixel\(const Vector3f &p_w, const SE3f &T_c_w\) const \{"),
    ],
    "keyframe_members.inc": [
        ("include/KeyFrame.h", r"inline Vector3f Pixel2Camera\(const Vector2f &p_p, float depth = 1\) const \{"),
    ],
    # out-of-class definitions (included at namespace ygz scope)
    "frame_bodies.inc": [
        ("src/Frame.cc", r"void Frame::AssignFeaturesToGrid\(\) \{"),
        ("src/Frame.cc", r"vector<size_t> Frame::GetFeaturesInArea\(const float &x, const float &y, const float &r, const int minLevel,"),
        ("src/Frame.cc", r"bool Frame::PosInGrid\(const cv::KeyPoint &kp, int &posX, int &posY\) \{"),
        ("src/Frame.cc", r"void Frame::ComputeStereoMatches\(\) \{"),
    ],
    "keyframe_bodies.inc": [
        ("src/KeyFrame.cc", r"vector<size_t> KeyFrame::GetFeaturesInArea\(const float &x, const float &y, const float &r\) const \{"),
        ("src/KeyFrame.cc", r"bool KeyFrame::IsInImage\(const float &x, const float &y\) const \{"),
        ("src/KeyFrame.cc", r"SE3f KeyFrame::GetPose\(\) \{"),
        ("src/KeyFrame.cc", r"Vector3f KeyFrame::GetCameraCenter\(\) \{"),
        ("src/KeyFrame.cc", r"Matrix3f KeyFrame::GetRotation\(\) \{"),
        ("src/KeyFrame.cc", r"Vector3f KeyFrame::GetTranslation\(\) \{"),
    ],
    "mappoint_bodies.inc": [
        ("src/MapPoint.cc", r"Vector3f MapPoint::GetWorldPos\(\) \{"),
        ("src/MapPoint.cc", r"Vector3f MapPoint::GetNormal\(\) \{"),
        ("src/MapPoint.cc", r"map<KeyFrame \*, size_t> MapPoint::GetObservations\(\) \{"),
        ("src/MapPoint.cc", r"int MapPoint::Observations\(\) \{"),
        ("src/MapPoint.cc", r"bool MapPoint::isBad\(\) \{"),
        ("src/MapPoint.cc", r"void MapPoint::ComputeDistinctiveDescriptors\(\) \{"),
        ("src/MapPoint.cc", r"cv::Mat MapPoint::GetDescriptor\(\) \{"),
        ("src/MapPoint.cc", r"int MapPoint::GetIndexInKeyFrame\(KeyFrame \*pKF\) \{"),
        ("src/MapPoint.cc", r"bool MapPoint::IsInKeyFrame\(KeyFrame \*pKF\) \{"),
        ("src/MapPoint.cc", r"float MapPoint::GetMinDistanceInvariance\(\) \{"),
        ("src/MapPoint.cc", r"float MapPoint::GetMaxDistanceInvariance\(\) \{"),
        ("src/MapPoint.cc", r"int MapPoint::PredictScale\(const float &currentDist, KeyFrame \*pKF\) \{"),
        ("src/MapPoint.cc", r"int MapPoint::PredictScale\(const float &currentDist, Frame \*pF\) \{"),
    ],
}


def cut(lines, first):
    """Lines of the definition starting at index `first`, up to the brace that closes its body."""
    depth, seen, out = 0, False, []
    for ln in lines[first:]:
        out.append(ln)
        code = re.sub(r"//.*", "", ln)
        code = re.sub(r'"(\\.|[^"\\])*"', '""', code)
        for ch in code:
            if ch == "{":
                depth += 1
                seen = True
            elif ch == "}":
                depth -= 1
        if seen and depth == 0:
            return out
    raise SystemExit("unbalanced braces")


def main():
    os.makedirs(OUT, exist_ok=True)
    cache = {}
    for name, items in TABLE.items():
        chunks = []
        for rel, pat in items:
            path = os.path.join(REF, rel)
            if path not in cache:
                cache[path] = open(path, encoding="utf-8", errors="replace").read().split("\n")
            lines = cache[path]
            hits = [i for i, ln in enumerate(lines) if re.search(pat, ln)]
            if len(hits) != 1:
                raise SystemExit("%s: %d definitions match %r" % (rel, len(hits), pat))
            body = cut(lines, hits[0])
            chunks.append("// ---- %s:%d-%d (copied at build time, unmodified)\n#line %d \"%s\"\n%s\n"
                          % (rel, hits[0] + 1, hits[0] + len(body), hits[0] + 1, path, "\n".join(body)))
        with open(os.path.join(OUT, name), "w") as f:
            f.write("// GENERATED by oracle/extract_ref_bodies.py from the reference tree at build time. Not tracked.\n")
            f.write("\n".join(chunks))
    print("extracted %d definitions into %s" % (sum(len(v) for v in TABLE.values()), OUT))


if __name__ == "__main__":
    main()
