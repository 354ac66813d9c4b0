#ifndef HUANG_FUNCTIONAL_H
#define HUANG_FUNCTIONAL_H
#include "AdaptationFunctional.h"
#endif
