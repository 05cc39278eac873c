// Dense Gaussian, float64: compiled kernel shapes.  Part 3 of 4: the shape list is split over
// several translation units only so that they compile in parallel.
#include "launch.cuh"
#define S(UNITM, EPL, LPC) HMCB_SHAPE(GaussianModel, HMCB_MODEL_GAUSSIAN, double, HMCB_F64, UNITM, EPL, LPC)
S(1, 26, 2) S(1, 14, 4) S(1, 8, 32) S(0, 4, 1)
