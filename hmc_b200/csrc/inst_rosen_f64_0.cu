// Pairwise Rosenbrock, float64: compiled kernel shapes.  Part 1 of 4: the shape list is split over
// several translation units only so that they compile in parallel.
#include "launch.cuh"
#define S(UNITM, EPL, LPC) HMCB_SHAPE(RosenbrockModel, HMCB_MODEL_ROSENBROCK, double, HMCB_F64, UNITM, EPL, LPC)
S(1, 2, 1) S(1, 52, 2) S(1, 14, 4) S(1, 8, 8) S(1, 4, 32)
