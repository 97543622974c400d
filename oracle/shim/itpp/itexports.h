/* Hand-written stand-in for IT++'s generated itexports.h (static build, no
 * symbol decoration).  Test infrastructure only. */
#ifndef ITEXPORTS_H
#define ITEXPORTS_H
#ifndef ITPP_EXPORT_TEMPLATE
#define ITPP_EXPORT_TEMPLATE extern
#endif
#ifndef ITPP_EXPORT
#define ITPP_EXPORT
#endif
#endif
