package work.pollochang.compression.image.core;

import work.pollochang.compression.image.learn.LearnedParams;
import work.pollochang.compression.image.learn.jpg.SimilarityKey;
import work.pollochang.compression.image.nativeffi.Icb;
import work.pollochang.compression.image.report.CompressionParams;

import java.awt.image.BufferedImage;
import java.io.IOException;
import java.io.OutputStream;
import java.nio.file.Files;
import java.nio.file.Path;
import java.util.Map;

import static work.pollochang.compression.image.tools.CacheTools.createKey;

/**
 * Drop-in replacement of the reference class of the same name (same package, same static
 * signatures, same boolean / exception contract), with the JPG target-size loop running on the
 * GPU through {@link Icb}.  NOT COMPILED IN THE BUILD IMAGE.
 *
 * To install: rename the reference's class to {@code ImageCompressionJpgStock} (it stays the
 * path for images the GPU library does not take: gray, CMYK, alpha, custom rasters) and add this
 * file.  Callers ({@code ImageCompression.compressImageIteratively}) do not change.
 *
 * What moved into the library (include/imgcomp_b200.h icb_jpg_fit): the cached try, the 0.85
 * scale loop, the float32 bisection with one full encode per step and the final save encode.
 * What stays here: the cache key, cache.get / cache.put and the file write -- the library hands
 * back the JFIF bytes and the learned (quality, scale).
 */
public final class ImageCompressionJpg {
    private ImageCompressionJpg() { }

    public static boolean compressJpgWithTargetSize(BufferedImage originalImage, long originalSize, Path outputFile,
                                                    CompressionParams params,
                                                    Map<SimilarityKey, LearnedParams> cache) throws IOException {
        SimilarityKey key = createKey(originalImage, originalSize);
        LearnedParams cached = cache.get(key);
        Icb.Fit fit = Icb.jpgFit(originalImage, params.quality(), params.targetMaxSizeBytes(),
                cached != null, cached != null ? cached.quality() : 0f, cached != null ? cached.scale() : 1.0);
        switch (fit.status()) {
            case Icb.OK -> {
                Files.write(outputFile, fit.jfif());
                if (!fit.usedCache()) cache.put(key, new LearnedParams(fit.bestQuality(), fit.scale()));
                return true;
            }
            case Icb.NO_FIT -> {
                return false;                                   // caller deletes the output: FAILED_COMPRESSION
            }
            case Icb.ERR_UNSUPPORTED -> {
                return ImageCompressionJpgStock.compressJpgWithTargetSize(originalImage, originalSize, outputFile, params, cache);
            }
            case Icb.ERR_OOM -> throw new OutOfMemoryError("icb_jpg_fit");          // FAILED_OUT_OF_MEMORY
            case Icb.ERR_INVALID -> throw new IllegalArgumentException("icb_jpg_fit: invalid argument");
            default -> throw new IOException("icb_jpg_fit status " + fit.status()); // FAILED_IO_ERROR
        }
    }

    /** The reference keeps this private; it is package-visible here so the golden dump can reach it. */
    static void compressJpgToStream(BufferedImage image, OutputStream os, float quality) throws IOException {
        if (quality < 0f || quality > 1f) throw new IllegalArgumentException("Quality out-of-bounds!");   // ImageWriteParam
        byte[] jfif = Icb.jpgEncode(image, quality);
        if (jfif == null) {
            ImageCompressionJpgStock.compressJpgToStreamStock(image, os, quality);
            return;
        }
        os.write(jfif);
    }
}
