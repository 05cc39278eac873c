// Pairwise Rosenbrock, float64: compiled kernel shapes.  Part 4 of 4: the shape list is split over
// several translation units only so that they compile in parallel.
#include "launch.cuh"
#define S(UNITM, EPL, LPC) HMCB_SHAPE(RosenbrockModel, HMCB_MODEL_ROSENBROCK, double, HMCB_F64, UNITM, EPL, LPC)
S(1, 20, 1) S(1, 28, 4) S(0, 14, 4) S(0, 28, 4) S(0, 8, 32)
