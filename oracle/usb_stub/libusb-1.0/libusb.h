/* TEST INFRASTRUCTURE: stand-in for <libusb-1.0/libusb.h>, which is not installed here.
 * The reference's software/ft601.h only needs these two opaque types to declare
 * struct ft601_context; no libusb function is called (oracle/ft601_bridge.c replaces
 * software/ft601.c). */
#ifndef ORACLE_LIBUSB_STUB_H
#define ORACLE_LIBUSB_STUB_H
#include <stdint.h>   /* the real header pulls this in; software/main.c relies on it */
typedef struct libusb_context libusb_context;
typedef struct libusb_device_handle libusb_device_handle;
#endif
