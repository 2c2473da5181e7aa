#include <mrpt/opengl/CSetOfObjects.h>
