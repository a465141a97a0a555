package work.pollochang.compression.image.core;

import work.pollochang.compression.image.learn.LearnedParams;
import work.pollochang.compression.image.learn.jpg.SimilarityKey;
import work.pollochang.compression.image.nativeffi.Icb;
import work.pollochang.compression.image.report.CompressionParams;

import java.io.IOException;
import java.nio.file.Files;
import java.nio.file.Path;
import java.util.Map;

/**
 * SURVEY section 8 row f2: the JPEG reader in front of the JPG seam, on the GPU.  NOT COMPILED IN THE BUILD IMAGE.
 *
 * The reference decodes every file on the host ({@code ImageCompression.decodeImageWithSubsampling},
 * core/ImageCompression.java:107-165) and hands the 36 MB raster of a 12 MP image to
 * {@code compressJpgWithTargetSize}.  For the files that make up its workload -- baseline 8-bit YCbCr 4:2:0 / 4:2:2 / 4:4:4, one
 * scan -- the library decodes the FILE on the GPU instead (include/imgcomp_b200.h
 * icb_jpg_file_fit): only the file crosses PCIe, the decode rule of :140-153 (source subsampling above 4096 px) is
 * applied on the device, and the fit is the same icb_jpg_fit.
 *
 * To install: in {@code ImageCompression.compressImage}, after {@code originalSize} is known and before the
 * {@code try (DecodedImage decodedImage = decodeImageWithSubsampling(...))} block (:65), add
 * <pre>
 *     if (isJpegName(inputPath)) {
 *         Boolean done = JpgFileFastPath.tryCompress(inputPath, originalSize, outputDir.resolve(inputPath.getFileName()), params, cache);
 *         if (done != null) { ... build the CompressionReport exactly as :80-95 do from `success` ... }
 *     }
 * </pre>
 * {@code null} means "not taken": the stock decode and the pixel seam run as before.
 */
public final class JpgFileFastPath {
    private JpgFileFastPath() { }

    /** @return TRUE / FALSE = what compressJpgWithTargetSize would have returned; null = use the stock path. */
    public static Boolean tryCompress(Path input, long fileSize, Path outputFile, CompressionParams params,
                                      Map<SimilarityKey, LearnedParams> cache) throws IOException {
        if (fileSize <= params.minSizeBytes()) return null;                 // :108-111: the stock path logs and skips
        byte[] file = Files.readAllBytes(input);
        int[] dims = Icb.jpgFileInfo(file);
        if (dims == null) return null;                                      // another kind of JPEG: ImageIO decodes it
        if (dims[0] <= params.minWidth() || dims[1] <= params.minHeight()) return null;   // :130-134 skip, stock path logs it
        // :140-153 the decode rule, to know the dimensions CacheTools.createKey (tools/CacheTools.java:14-21) will see
        int maxDim = Math.max(dims[0], dims[1]), s = 1;
        if (maxDim > 4096) s = Integer.highestOneBit(Math.max(1, (int) Math.floor((double) maxDim / 4096)));
        int w = (dims[0] + s - 1) / s, h = (dims[1] + s - 1) / s;
        SimilarityKey key = new SimilarityKey(w / 100, h / 100, fileSize / 102400);
        LearnedParams cached = cache.get(key);
        Icb.FileFit ff = Icb.jpgFileFit(file, params.quality(), params.targetMaxSizeBytes(),
                cached != null, cached != null ? cached.quality() : 0f, cached != null ? cached.scale() : 1.0);
        Icb.Fit fit = ff.fit();
        switch (fit.status()) {
            case Icb.OK -> {
                Files.write(outputFile, fit.jfif());                        // core/ImageCompressionJpg.java:258 / :229
                if (!fit.usedCache()) cache.put(key, new LearnedParams(fit.bestQuality(), fit.scale()));   // :111
                return Boolean.TRUE;
            }
            case Icb.NO_FIT -> {
                return Boolean.FALSE;                                       // :120-121
            }
            case Icb.ERR_UNSUPPORTED, Icb.ERR_INVALID -> {
                return null;                                                // stream the parallel decoder gave back / broken file
            }
            case Icb.ERR_OOM -> throw new OutOfMemoryError("icb_jpg_file_fit");
            default -> throw new IOException("icb_jpg_file_fit status " + fit.status());
        }
    }
}
