// TEST INFRASTRUCTURE (oracle build only): xmlSAXUserParseFile() over expat.
// See shim/libxml/parser.h. SAX1 semantics reproduced: startElement gets the
// element name and a NULL-terminated array of alternating attribute key/value
// strings (NULL itself when the element has no attributes, as libxml2 does).
#include <libxml/parser.h>
#include <expat.h>
#include <cstdio>
#include <vector>

namespace {
struct Ctx { xmlSAXHandler *sax; void *user; };
void XMLCALL onStart(void *ud, const XML_Char *name, const XML_Char **atts)
{
	Ctx *c = (Ctx *)ud;
	if(!c->sax->startElement) return;
	const xmlChar **a = (atts && atts[0]) ? (const xmlChar **)atts : nullptr;
	c->sax->startElement(c->user, (const xmlChar *)name, a);
}
void XMLCALL onEnd(void *ud, const XML_Char *name)
{
	Ctx *c = (Ctx *)ud;
	if(c->sax->endElement) c->sax->endElement(c->user, (const xmlChar *)name);
}
}

extern "C" int xmlSAXUserParseFile(xmlSAXHandler *sax, void *user_data, const char *filename)
{
	FILE *f = std::fopen(filename, "rb");
	if(!f)
	{
		if(sax->fatalError) sax->fatalError(user_data, "cannot open file %s", filename);
		return -1;
	}
	XML_Parser p = XML_ParserCreate(nullptr);
	Ctx ctx{sax, user_data};
	XML_SetUserData(p, &ctx);
	XML_SetElementHandler(p, onStart, onEnd);
	if(sax->startDocument) sax->startDocument(user_data);
	std::vector<char> buf(1 << 20);
	int rc = 0;
	for(;;)
	{
		size_t n = std::fread(buf.data(), 1, buf.size(), f);
		int done = n < buf.size();
		if(XML_Parse(p, buf.data(), (int)n, done) == XML_STATUS_ERROR)
		{
			if(sax->fatalError) sax->fatalError(user_data, "%s at line %lu", XML_ErrorString(XML_GetErrorCode(p)), (unsigned long)XML_GetCurrentLineNumber(p));
			rc = -1;
			break;
		}
		if(done) break;
	}
	if(sax->endDocument) sax->endDocument(user_data);
	XML_ParserFree(p);
	std::fclose(f);
	return rc;
}
