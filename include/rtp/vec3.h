// vec3.h -- same name as the reference header (src/vec3.h); the API lives in scene_api.h
#include "scene_api.h"
