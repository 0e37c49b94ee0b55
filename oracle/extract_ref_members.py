#!/usr/bin/env python3
"""TEST INFRASTRUCTURE ONLY — run by oracle/Makefile at build time, never imported by the product.

Writes oracle/_ref/gen/ref_members.cc: the UNMODIFIED bodies of the Frame / KeyFrame / MapPoint member functions that
/root/reference/src/ORBmatcher.cc calls (and Frame::ComputeStereoMatches, MapPoint::ComputeDistinctiveDescriptors, which
are on the hot path themselves), cut out of the reference sources where they lie. The whole translation units cannot be
compiled here (boost serialization, Eigen, g2o, fbow vocabulary I/O, Map, KeyFrameDatabase ... none installed), so each
definition is located by its qualified name at column 0 and copied verbatim up to its closing brace at column 0; the
class declarations they compile against are oracle/shim_matcher/ref_slam_types.h. The output lands under oracle/_ref/
(git-ignored): no reference source text is committed.

usage: extract_ref_members.py <reference root> <output .cc>
"""
import re
import sys

# (file, qualified name, substring that must occur in the signature line (overload selector) or None)
WANT = [
    ("src/Frame.cc", "Frame::AssignFeaturesToGrid", None),
    ("src/Frame.cc", "Frame::GetFeaturesInArea", None),
    ("src/Frame.cc", "Frame::PosInGrid", None),
    ("src/Frame.cc", "Frame::ComputeStereoMatches", None),
    ("src/KeyFrame.cc", "KeyFrame::GetCameraCenter", None),
    ("src/KeyFrame.cc", "KeyFrame::GetRotation", None),
    ("src/KeyFrame.cc", "KeyFrame::GetTranslation", None),
    ("src/KeyFrame.cc", "KeyFrame::AddMapPoint", None),
    ("src/KeyFrame.cc", "KeyFrame::EraseMapPointMatch", "const size_t"),
    ("src/KeyFrame.cc", "KeyFrame::EraseMapPointMatch", "MapPoint*"),
    ("src/KeyFrame.cc", "KeyFrame::ReplaceMapPointMatch", None),
    ("src/KeyFrame.cc", "KeyFrame::GetMapPoints", None),
    ("src/KeyFrame.cc", "KeyFrame::GetMapPointMatches", None),
    ("src/KeyFrame.cc", "KeyFrame::GetMapPoint", None),
    ("src/KeyFrame.cc", "KeyFrame::GetFeaturesInArea", None),
    ("src/KeyFrame.cc", "KeyFrame::IsInImage", None),
    ("src/MapPoint.cc", "MapPoint::GetWorldPos", None),
    ("src/MapPoint.cc", "MapPoint::GetNormal", None),
    ("src/MapPoint.cc", "MapPoint::AddObservation", None),
    ("src/MapPoint.cc", "MapPoint::EraseObservation", None),
    ("src/MapPoint.cc", "MapPoint::GetObservations", None),
    ("src/MapPoint.cc", "MapPoint::Observations", None),
    ("src/MapPoint.cc", "MapPoint::SetBadFlag", None),
    ("src/MapPoint.cc", "MapPoint::Replace", None),
    ("src/MapPoint.cc", "MapPoint::isBad", None),
    ("src/MapPoint.cc", "MapPoint::IncreaseVisible", None),
    ("src/MapPoint.cc", "MapPoint::IncreaseFound", None),
    ("src/MapPoint.cc", "MapPoint::ComputeDistinctiveDescriptors", None),
    ("src/MapPoint.cc", "MapPoint::GetDescriptor", None),
    ("src/MapPoint.cc", "MapPoint::GetIndexInKeyFrame", None),
    ("src/MapPoint.cc", "MapPoint::IsInKeyFrame", None),
    ("src/MapPoint.cc", "MapPoint::GetMinDistanceInvariance", None),
    ("src/MapPoint.cc", "MapPoint::GetMaxDistanceInvariance", None),
    ("src/MapPoint.cc", "MapPoint::PredictScale", "KeyFrame*"),
    ("src/MapPoint.cc", "MapPoint::PredictScale", ",Frame*"),
]
# static data members, copied as whole lines
STATICS = [
    ("src/Frame.cc", re.compile(r"^float Frame::")),
    ("src/MapPoint.cc", re.compile(r"^mutex MapPoint::mGlobalMutex")),
]


def cut(lines, qname, selector, path):
    pat = re.compile(r"^[A-Za-z_][\w:<>\*&, ]*\b" + re.escape(qname) + r"\(")
    for i, ln in enumerate(lines):
        if pat.match(ln) and (selector is None or selector.replace(" ", "") in ln.replace(" ", "")):
            for j in range(i + 1, len(lines)):
                if lines[j].startswith("}"):
                    return i, j
            break
    raise SystemExit(f"extract_ref_members: {qname} ({selector}) not found in {path}")


def main(ref, out):
    cache = {}
    chunks = []
    for path, rx in STATICS:
        lines = cache.setdefault(path, open(f"{ref}/{path}").read().split("\n"))
        for i, ln in enumerate(lines):
            if rx.match(ln):
                chunks.append(f"// {path}:{i + 1}\n{ln}\n")
    for path, qname, sel in WANT:
        lines = cache.setdefault(path, open(f"{ref}/{path}").read().split("\n"))
        i, j = cut(lines, qname, sel, path)
        chunks.append(f"// {path}:{i + 1}-{j + 1}\n" + "\n".join(lines[i:j + 1]) + "\n")
    with open(out, "w") as f:
        f.write("// GENERATED at build time by oracle/extract_ref_members.py from the reference tree — do not commit.\n")
        f.write('#include "ORBmatcher.h"\n#include <thread>\nusing namespace std;\nnamespace ORB_SLAM2 {\n\n')
        f.write("\n".join(chunks))
        f.write("\n}  // namespace ORB_SLAM2\n")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
