// TEST INFRASTRUCTURE (oracle build only).
// Minimal stand-in for the part of libxml2's SAX1 interface that the reference's
// XML importer uses (/root/reference/src/common/import_xml.cc:31,69-149,166):
// xmlChar, xmlSAXHandler (positional initialisation: slots 12..15 are
// startDocument/endDocument/startElement/endElement, 21..23 warning/error/
// fatalError) and xmlSAXUserParseFile(). libxml2 is an un-vendored third-party
// dependency that is absent from this image; the implementation in
// xml_expat_shim.cc drives the same callbacks from expat.
#pragma once
#include <stdarg.h>
#include <stdio.h>
#ifdef __cplusplus
extern "C" {
#endif
typedef unsigned char xmlChar;
typedef void (*yafshim_void_fn)(void *);
typedef void (*yafshim_start_el_fn)(void *, const xmlChar *, const xmlChar **);
typedef void (*yafshim_end_el_fn)(void *, const xmlChar *);
typedef void (*yafshim_msg_fn)(void *, const char *, ...);
typedef struct _xmlSAXHandler
{
	void *slot0_11[12];
	yafshim_void_fn startDocument;      // 12
	yafshim_void_fn endDocument;        // 13
	yafshim_start_el_fn startElement;   // 14
	yafshim_end_el_fn endElement;       // 15
	void *slot16;                       // reference
	void *characters;                   // 17
	void *slot18_20[3];
	yafshim_msg_fn warning;             // 21
	yafshim_msg_fn error;               // 22
	yafshim_msg_fn fatalError;          // 23
	void *tail[8];
} xmlSAXHandler;
int xmlSAXUserParseFile(xmlSAXHandler *sax, void *user_data, const char *filename);
#ifdef __cplusplus
}
#endif
