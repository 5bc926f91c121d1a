// "ERROR:" / "WARN:" / "INFO:" tags of src/include/sens_loc/util/console.h:45-67:
// coloured on a terminal, plain otherwise (as rang behaves).
#ifndef DCB200_DEPTH2X_CONSOLE_H
#define DCB200_DEPTH2X_CONSOLE_H

#include <iostream>
#include <unistd.h>

namespace sens_loc::apps::console {
inline bool tty() {
    static const bool t = ::isatty(2) != 0;
    return t;
}
inline const char* bold() { return tty() ? "\033[1m" : ""; }
inline const char* reset() { return tty() ? "\033[0m" : ""; }
inline const char* err() { return tty() ? "\033[31m\033[1mERROR:\033[0m " : "ERROR: "; }
inline const char* warn() { return tty() ? "\033[33m\033[1mWARN:\033[0m " : "WARN: "; }
inline const char* info() { return tty() ? "\033[1mINFO:\033[0m " : "INFO: "; }
}  // namespace sens_loc::apps::console
#endif
