// color.h -- same name as the reference header (src/color.h); the API lives in scene_api.h
#include "scene_api.h"
