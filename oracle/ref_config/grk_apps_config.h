/* hand-written stand-in for the cmake-generated grk_apps_config.h of the reference applications:
 * no optional image libraries (test infrastructure, see oracle/Makefile) */
#include "grk_config_private.h"
