// Pairwise Rosenbrock, float64: compiled kernel shapes.  Part 2 of 4: the shape list is split over
// several translation units only so that they compile in parallel.
#include "launch.cuh"
#define S(UNITM, EPL, LPC) HMCB_SHAPE(RosenbrockModel, HMCB_MODEL_ROSENBROCK, double, HMCB_F64, UNITM, EPL, LPC)
S(1, 4, 1) S(1, 50, 1) S(1, 16, 8) S(0, 4, 1) S(0, 10, 1)
