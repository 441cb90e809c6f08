// TEST INFRASTRUCTURE ONLY. Empty shadow of the reference's utils/eigen_util.hh
// (unused helpers that need the real Eigen).
