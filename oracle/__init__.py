"""TEST INFRASTRUCTURE - never imported by the product (stereo_vo_b200/); only tests/,
__graft_entry__.smoke() / build() and bench.py's CPU-baseline legs use it.

  cv2_frontend.py          the reference restated call for call on Python cv2 (the oracle the
                           CUDA path is compared with); pinned by ref_frontend / ref_ba below
  restate.py               numpy restatement of the arithmetic inside those OpenCV calls
  reprojection_factor.py   ReprojectionFactor::Evaluate in numpy
  build_ref.py             compiles the reference's OWN sources where they lie into _ref/
                           (git-ignored): libref_frontend.so (image_processor.cpp,
                           feature_tracker.cpp, bundle_adjuster.cpp, reprojection_factor.cpp; the
                           OpenCV algorithms served by cv2 through callbacks) and libref_ba.so
  ref_stubs/               the stand-in Eigen / OpenCV / Ceres headers those builds use
  ref_harness.cpp, ref_frontend_harness.cpp   C entry points over the reference's classes
  ref_ba.py, ref_frontend.py                  ctypes drivers of the two libraries
"""
