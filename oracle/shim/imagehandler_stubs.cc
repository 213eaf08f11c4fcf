// TEST INFRASTRUCTURE (oracle build only).
// The reference's PNG/JPG/TIF/EXR handlers (/root/reference/src/imagehandler/
// imagehandler_{png,jpg,tif,exr}.cc) include libpng/libjpeg/libtiff/OpenEXR
// headers, none of which exist in this image. The dispatcher
// (/root/reference/src/imagehandler/imagehandler.cc:31-50) still references
// their factories, so they are defined here to refuse the format. TGA and HDR
// handlers are built from the reference's own sources. Nothing on the hot path
// (intersection, integrators, lights, materials, film) touches these.
#include "imagehandler/imagehandler_png.h"
#include "imagehandler/imagehandler_jpg.h"
#include "imagehandler/imagehandler_tif.h"
#include "imagehandler/imagehandler_exr.h"
#include "common/logging.h"

BEGIN_YAFARAY
#define YAF_STUB(cls, name) \
	ImageHandler *cls::factory(ParamMap &, RenderEnvironment &) \
	{ Y_ERROR << "oracle build: image format '" name "' not available" << YENDL; return nullptr; }
YAF_STUB(PngHandler, "png")
YAF_STUB(JpgHandler, "jpg")
YAF_STUB(TifHandler, "tif")
YAF_STUB(ExrHandler, "exr")
END_YAFARAY
