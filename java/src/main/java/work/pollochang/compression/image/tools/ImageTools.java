package work.pollochang.compression.image.tools;

import work.pollochang.compression.image.nativeffi.Icb;

import java.awt.Graphics2D;
import java.awt.RenderingHints;
import java.awt.image.BufferedImage;

/**
 * Drop-in replacement of the reference's ImageTools: same signature and result type, the bilinear
 * drawImage runs on the GPU for byte rasters (TYPE_3BYTE_BGR, TYPE_4BYTE_ABGR, TYPE_BYTE_GRAY --
 * what the ImageIO JPEG / PNG readers produce); every other raster type keeps Graphics2D.
 * NOT COMPILED IN THE BUILD IMAGE.
 */
public final class ImageTools {
    private ImageTools() { }

    public static BufferedImage resizeImage(BufferedImage originalImage, double scale) {
        int[] d = Icb.resizeDims(originalImage.getWidth(), originalImage.getHeight(), scale);
        int imageType = originalImage.getType();
        if (imageType == BufferedImage.TYPE_CUSTOM) {
            imageType = originalImage.getAlphaRaster() != null ? BufferedImage.TYPE_INT_ARGB : BufferedImage.TYPE_INT_RGB;
        }
        BufferedImage resized = new BufferedImage(d[0], d[1], imageType);
        if (imageType == originalImage.getType() && Icb.resizeInto(originalImage, resized)) return resized;
        Graphics2D g2d = resized.createGraphics();
        try {
            g2d.setRenderingHint(RenderingHints.KEY_INTERPOLATION, RenderingHints.VALUE_INTERPOLATION_BILINEAR);
            g2d.setRenderingHint(RenderingHints.KEY_RENDERING, RenderingHints.VALUE_RENDER_QUALITY);
            g2d.setRenderingHint(RenderingHints.KEY_ANTIALIASING, RenderingHints.VALUE_ANTIALIAS_ON);
            g2d.drawImage(originalImage, 0, 0, d[0], d[1], null);
        } finally {
            g2d.dispose();
        }
        return resized;
    }
}
