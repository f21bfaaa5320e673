// See j2k_cuda_codec.h.  Translates j2k::Buffer / FileInfo / InputFile / OutputFile to the plain-C calls of
// include/j2kcuda.h and throws j2k::Exception on a non-zero status, the way OpenJPEGCodec reports opj_*
// failures (src/common/j2k_openjpeg_codec.cpp:452-758).
#include "j2k_cuda_codec.h"
#include "j2k_exception.h"
#include "j2kcuda.h"

#include <vector>
#include <string>
#include <cstdlib>
#include <cstring>
#include <cstdio>

namespace j2k
{

// How much of CompressionSettings WriteFile honours.
//   AS_WRITTEN  exactly what OpenJPEGCodec::WriteFile passes on today (src/common/j2k_openjpeg_codec.cpp:703-719): the number
//               of layers, cp_disto_alloc and the tile size; everything else is the library default (lossless 5/3, no MCT,
//               6 resolutions, 64x64 code-blocks, LRCP, raw codestream).  Output is byte-identical to the reference codec's
//               (tests/test_integration_plugin.py).
//   FULL        (default) also what the reference marks "TODO: copy more settings": reversible / irreversible, size and
//               quality targets, progression order, the DCI profiles, RCT/ICT for RGB, JP2 / JPX wrapping.
// Chosen by CudaCodec::SetMode() or the environment variable J2K_CUDA_CODEC_MODE=as-written|full.
static int g_mode = -1;
void CudaCodec::SetMode(Mode m) { g_mode = m; }
CudaCodec::Mode CudaCodec::GetMode()
{
	if(g_mode < 0)
	{
		const char *e = getenv("J2K_CUDA_CODEC_MODE");
		g_mode = (e != NULL && strcmp(e, "as-written") == 0) ? AS_WRITTEN : FULL;
	}
	return static_cast<Mode>(g_mode);
}

static void LogToStdout(void *, int level, const char *msg)       // the reference prints library messages (j2k_openjpeg_codec.cpp:156-172)
{
	printf("libj2kcuda %s: %s\n", (level == 0 ? "error" : "info"), msg);
}

static void InitOnce()
{
	static bool done = false;
	if(!done)
	{
		j2kcuda_init(-1, LogToStdout, NULL);
		done = true;
	}
}

static std::vector<uint8_t> ReadAll(InputFile &file)
{
	const size_t n = file.FileSize();
	std::vector<uint8_t> bytes(n);
	file.Seek(0);
	if(n && file.Read(&bytes[0], n) != n)
		throw Exception("Error reading file.");
	return bytes;
}

static void Check(int status)
{
	if(status != J2KCUDA_OK)
		throw Exception(std::string("libj2kcuda: ") + j2kcuda_last_error());
}

static void ToPlanes(const Buffer &buffer, j2kcuda_plane planes[J2K_CODEC_MAX_CHANNELS])
{
	for(int i=0; i < buffer.channels; i++)
	{
		const Channel &c = buffer.channel[i];
		planes[i].buf = c.buf;
		planes[i].colbytes = c.colbytes;
		planes[i].rowbytes = c.rowbytes;
		planes[i].width = c.width;
		planes[i].height = c.height;
		planes[i].sample_type = (c.sampleType == UCHAR ? J2KCUDA_UCHAR : c.sampleType == USHORT ? J2KCUDA_USHORT :
									c.sampleType == UINT ? J2KCUDA_UINT : J2KCUDA_INT);
		planes[i].depth = c.depth;
		planes[i].sgnd = c.sgnd;
		planes[i].reserved = 0;
	}
}

void CudaCodec::GetFileInfo(InputFile &file, FileInfo &info)
{
	InitOnce();
	const std::vector<uint8_t> bytes = ReadAll(file);
	j2kcuda_info hdr;
	Check( j2kcuda_probe(bytes.empty() ? NULL : &bytes[0], bytes.size(), &hdr) );

	// same fields OpenJPEGCodec::GetFileInfo fills from opj_read_header_ex (j2k_openjpeg_codec.cpp:222-429)
	info.width = hdr.width;
	info.height = hdr.height;
	info.channels = hdr.ncomp;
	info.depth = hdr.prec[0];
	for(unsigned int i=0; i < hdr.ncomp && i < J2K_CODEC_MAX_CHANNELS; i++)
		info.subsampling[i] = Subsampling(hdr.dx[i], hdr.dy[i]);
	info.format = (hdr.is_jp2 ? JP2 : J2C);
	info.colorSpace = (hdr.ncomp >= 3 ? sRGB : sLUM);
	info.alpha = (hdr.ncomp == 2 || hdr.ncomp == 4 ? STRAIGHT : NO_ALPHA);

	if(hdr.is_jp2)
	{
		// the JP2 header boxes, as OpenJPEGCodec::GetFileInfo takes them from opj_read_header_ex (j2k_openjpeg_codec.cpp:303-408)
		j2kcuda_jp2_info jp2;
		Check( j2kcuda_probe_jp2(&bytes[0], bytes.size(), &jp2) );

		if(jp2.brand == 0x6a707820)               // 'jpx '
			info.format = JPX;

		enum { CS_CMYK = 12, CS_sRGB = 16, CS_GRAY = 17, CS_sYCC = 18, CS_esYCC = 19 };
		if(jp2.colr_method == 1)
			info.colorSpace = (jp2.enumcs == CS_CMYK ? CMYK : jp2.enumcs == CS_sRGB ? sRGB : jp2.enumcs == CS_GRAY ? sLUM :
								jp2.enumcs == CS_sYCC ? sYCC : jp2.enumcs == CS_esYCC ? esYCC : UNKNOWN_COLOR_SPACE);
		else if((jp2.colr_method == 2 || jp2.colr_method == 3) && jp2.icc_length > 0)
		{
			info.iccProfile = malloc(jp2.icc_length);      // freed by RGBAinputFile (j2k_rgba_file.cpp:50-54)
			if(info.iccProfile == NULL)
				throw Exception("out of memory");
			info.profileLen = jp2.icc_length;
			memcpy(info.iccProfile, &bytes[jp2.icc_offset], jp2.icc_length);
			info.colorSpace = (hdr.ncomp >= 3 ? iccRGB : hdr.ncomp == 1 ? iccLUM : iccANY);
		}

		if(jp2.cdef_n)
		{
			info.alpha = NO_ALPHA;
			for(unsigned int i=0; i < jp2.cdef_n && i < J2K_CODEC_MAX_CHANNELS; i++)
			{
				if(jp2.cdef_cn[i] >= J2K_CODEC_MAX_CHANNELS)
					continue;
				// colour asoc 1..3 -> R, G, B; opacity (typ 1 straight, typ 2 premultiplied) -> ALPHA
				if(jp2.cdef_typ[i] == 0 && jp2.cdef_asoc[i] >= 1 && jp2.cdef_asoc[i] <= 3)
					info.channelMap[jp2.cdef_cn[i]] = (jp2.cdef_asoc[i] == 1 ? RED : jp2.cdef_asoc[i] == 2 ? GREEN : BLUE);
				else if(jp2.cdef_typ[i] == 1 || jp2.cdef_typ[i] == 2)
				{
					info.channelMap[jp2.cdef_cn[i]] = ALPHA;
					info.alpha = (jp2.cdef_typ[i] == 2 ? PREMULTIPLIED : STRAIGHT);
				}
			}
		}

		if(jp2.pclr_entries && jp2.pclr_entries <= J2K_CODEC_MAX_LUT_ENTRIES && jp2.pclr_channels <= J2K_CODEC_MAX_CHANNELS)
		{
			std::vector<uint32_t> entries(jp2.pclr_entries * jp2.pclr_channels);
			if(j2kcuda_jp2_palette(&bytes[0], bytes.size(), &entries[0], entries.size()) < 0)
				throw Exception(j2kcuda_last_error());
			info.LUTsize = jp2.pclr_entries;
			for(unsigned int e=0; e < jp2.pclr_entries; e++)
				for(unsigned int c=0; c < jp2.pclr_channels; c++)
					info.LUT[e].channel[c] = static_cast<unsigned char>(entries[e * jp2.pclr_channels + c]);
			for(unsigned int i=0; i < jp2.cmap_n && i < J2K_CODEC_MAX_CHANNELS; i++)
				if(jp2.cmap_mtyp[i] == 1 && jp2.cmap_pcol[i] < J2K_CODEC_MAX_CHANNELS)
					info.LUTmap[i] = info.channelMap[jp2.cmap_pcol[i]];
		}
	}

	info.settings.reversible = hdr.reversible;
	info.settings.layers = hdr.numlayers;
	info.settings.tileSize = (hdr.tiles_x * hdr.tiles_y > 1 ? hdr.tile_w : 0);
}

void CudaCodec::ReadFile(InputFile &file, const Buffer &buffer, unsigned int subsample, Progress *progress)
{
	InitOnce();
	const std::vector<uint8_t> bytes = ReadAll(file);
	unsigned int reduce = 0;                     // cp_reduce = log2(subsample), j2k_openjpeg_codec.cpp:496-501
	while((1u << (reduce + 1)) <= subsample)
		reduce++;

	j2kcuda_plane planes[J2K_CODEC_MAX_CHANNELS];
	ToPlanes(buffer, planes);

	Check( j2kcuda_decode(bytes.empty() ? NULL : &bytes[0], bytes.size(), reduce, planes, buffer.channels, NULL, NULL) );

	if(progress && progress->abortProc && progress->abortProc(progress->refCon))   // polled once, as the reference does (:515-528)
		progress->keepGoing = false;
}

void CudaCodec::WriteFile(OutputFile &file, const FileInfo &info, const Buffer &buffer, Progress *progress)
{
	InitOnce();
	const bool asWritten = (GetMode() == AS_WRITTEN);

	j2kcuda_params p;
	j2kcuda_default_params(&p);
	p.width = info.width;
	p.height = info.height;
	p.ncomp = buffer.channels;
	p.prec = info.depth;
	p.sgnd = 0;

	const CompressionSettings &s = info.settings;
	const unsigned int maxLayers = sizeof(p.rates) / sizeof(p.rates[0]);
	const unsigned int layers = (s.layers < 1 ? 1 : s.layers > maxLayers ? maxLayers : s.layers);
	if(s.tileSize && (s.tileSize < info.width || s.tileSize < info.height))
		p.tile_w = p.tile_h = s.tileSize;

	if(asWritten)
	{
		p.mct = 0;                               // library default tcp_mct (openjpeg.cpp:706-728); the reference never sets it
		p.numlayers = layers;                    // params.tcp_numlayers = info.settings.layers, all rates 0: lossless layers
	}
	else
	{
		p.mct = (buffer.channels >= 3 ? 1 : 0);  // what opj_compress does for RGB (bin/jp2/opj_compress.cpp:1399-1400)
		p.numlayers = (s.method == LOSSLESS ? 1 : layers);
		p.prog_order = (s.order == LRCP ? J2KCUDA_LRCP : s.order == RLCP ? J2KCUDA_RLCP : s.order == RPCL ? J2KCUDA_RPCL :
						s.order == PCRL ? J2KCUDA_PCRL : J2KCUDA_CPRL);

		if(s.method == CINEMA)
		{
			p.profile = (s.dciProfile == DCI_4K ? J2KCUDA_PROFILE_CINEMA_4K : J2KCUDA_PROFILE_CINEMA_2K);
			p.max_cs_size = s.fileSize * 1024;   // the caller pre-computes KiB per frame (src/aftereffects/j2k.cpp:817-830); 0 = profile default
		}
		else if(s.method == SIZE)
		{
			if(s.fileSize == 0)
				throw Exception("Target file size is zero.");
			p.irreversible = !s.reversible;
			const double raw = (double)info.width * info.height * buffer.channels * info.depth / 8.0;
			const double ratio = raw / ((double)s.fileSize * 1024.0);
			p.rates[p.numlayers - 1] = (ratio > 1.0 ? ratio : 1.0);
			for(int i = (int)p.numlayers - 2; i >= 0; i--)
				p.rates[i] = p.rates[i + 1] * 2.0;
		}
		else if(s.method == QUALITY)
		{
			p.irreversible = !s.reversible;
			const int q = (s.quality > 100 ? 100 : s.quality);
			p.rates[p.numlayers - 1] = 1.0 + (100 - q) * 0.5;   // quality slider -> compression ratio (the reference has no mapping: "TODO")
			for(int i = (int)p.numlayers - 2; i >= 0; i--)
				p.rates[i] = p.rates[i + 1] * 2.0;
		}
	}

	j2kcuda_plane planes[J2K_CODEC_MAX_CHANNELS];
	ToPlanes(buffer, planes);

	uint8_t *out = NULL;
	size_t out_len = 0;
	Check( j2kcuda_encode(&p, planes, buffer.channels, &out, &out_len) );

	if(!asWritten && (info.format == JP2 || info.format == JPX))
	{
		// JP2 boxes around the codestream (the OpenJPEG path of the reference writes raw J2K only: jp2 is a "TODO" there,
		// j2k_openjpeg_codec.cpp:613-614; the boxes are those opj_jp2_setup_encoder would produce)
		j2kcuda_jp2_params w;
		memset(&w, 0, sizeof(w));
		w.width = info.width;
		w.height = info.height;
		w.ncomp = buffer.channels;
		for(int i=0; i < buffer.channels; i++)
		{
			w.prec[i] = info.depth;
			w.sgnd[i] = 0;
		}
		w.enumcs = (buffer.channels >= 3 ? 16 : 17);
		if(info.iccProfile != NULL && info.profileLen > 0 && !IssRGBProfile(info.iccProfile, info.profileLen))
		{
			w.icc = static_cast<const uint8_t *>(info.iccProfile);
			w.icc_len = info.profileLen;
		}
		w.alpha_comp = (info.alpha != NO_ALPHA && (buffer.channels == 2 || buffer.channels == 4) ? buffer.channels - 1 : -1);

		uint8_t *boxed = NULL;
		size_t boxed_len = 0;
		const int rc = j2kcuda_wrap_jp2(out, out_len, &w, &boxed, &boxed_len);
		j2kcuda_free(out);
		Check(rc);
		out = boxed;
		out_len = boxed_len;
	}

	const size_t written = file.Write(out, out_len);      // one contiguous file image, one Write
	j2kcuda_free(out);

	if(written != out_len)
		throw Exception("Error writing file.");

	if(progress && progress->abortProc && progress->abortProc(progress->refCon))
		progress->keepGoing = false;
}

}; // namespace j2k
