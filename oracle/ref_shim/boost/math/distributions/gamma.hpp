// TEST INFRASTRUCTURE: QMOM.C:30 includes this Boost header but uses nothing from it; Boost is
// absent from this image.  Intentionally empty.
