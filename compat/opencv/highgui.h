#include "cv.h"
