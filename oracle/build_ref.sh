#!/bin/bash
# TEST INFRASTRUCTURE.  Compiles the reference's own hot-path sources, from where they lie under
# /root/reference, against oracle/refshim (a stand-in for the OpenCV/Eigen headers that are not
# installed here; its primitives come from oracle/cv_prims.hpp) into oracle/_ref/libref_orb.so.
# Flags as the reference's Release build (-O3 -DNDEBUG would drop its asserts; keep them on).
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${REFERENCE_ROOT:-/root/reference}"
[ -f "$REF/Features/ORBextractor.cpp" ] || { echo "build_ref.sh: no reference checkout at $REF"; exit 0; }
mkdir -p "$HERE/_ref"
# Initializer.cpp needs calib3d and Eigen as a whole; only its constructor and the two scoring functions
# (CheckHomography, CheckFundamental: plain float arithmetic) are cut out, each from its signature line to the
# first closing brace in column 0, into a generated include that stays under _ref/.
awk '/^Initializer::Initializer\(/ || /^float Initializer::CheckHomography\(/ || /^float Initializer::CheckFundamental\(/ {on=1}
     on {print} on && /^}/ {on=0}' "$REF/Initialization/Initializer.cpp" > "$HERE/_ref/initializer_slice.inc"
# ComputeKeyPointsOld (Features/ORBextractor.cpp:1207-1496) is compiled into the library above but never
# called: operator() calls ComputeKeyPointsOctTree (:1552) and keeps the old policy as a commented-out line (:1555).  To run
# the reference's own body, a copy of ORBextractor.cpp with that ONE call swapped is generated into _ref/ (git-ignored, never
# committed) and compiled into the same library under another namespace name.
sed 's|^  ComputeKeyPointsOctTree(allKeypoints);$|  ComputeKeyPointsOld(allKeypoints);|' "$REF/Features/ORBextractor.cpp" > "$HERE/_ref/ORBextractor_oldpolicy.cpp"
grep -q '^  ComputeKeyPointsOld(allKeypoints);$' "$HERE/_ref/ORBextractor_oldpolicy.cpp" || { echo "build_ref.sh: call site not found"; exit 1; }
OLD="-DORB_SLAM_Tracking=ORB_SLAM_Tracking_OldPolicy"
g++ -O3 -std=c++17 -fPIC -c -ffp-contract=off -w $OLD -I"$HERE/refshim" -I"$REF" "$HERE/_ref/ORBextractor_oldpolicy.cpp" -o "$HERE/_ref/oldpolicy.o"
g++ -O3 -std=c++17 -fPIC -c -ffp-contract=off -w $OLD -I"$HERE/refshim" -I"$REF" "$HERE/ref_old_driver.cpp" -o "$HERE/_ref/oldpolicy_driver.o"
g++ -O3 -std=c++17 -fPIC -shared -pthread -ffp-contract=off -w \
  -I"$HERE/refshim" -I"$REF" -I"$REF/Thirdparty/DBoW2/include" -I"$REF/Thirdparty/DBoW2/include/DBoW2" \
  "$REF/Features/ORBextractor.cpp" "$REF/Features/ORBmatcher.cpp" "$REF/SlamTypes/Frame.cpp" \
  "$REF/Thirdparty/DBoW2/src/FORB.cpp" "$REF/Thirdparty/DBoW2/src/BowVector.cpp" "$REF/Thirdparty/DBoW2/src/FeatureVector.cpp" \
  "$HERE/ref_driver.cpp" "$HERE/ref_init_driver.cpp" "$HERE/_ref/oldpolicy.o" "$HERE/_ref/oldpolicy_driver.o" -I"$HERE/_ref" -o "$HERE/_ref/libref_orb.so"
echo "built $HERE/_ref/libref_orb.so"
