-- | Drop-in for the reference's @Graphics.Ray.Texture@ (Texture.hs:2-5): same exported names and argument meaning,
-- but a texture is DATA (what the C ABI can carry) instead of a closure.  'solidTexture', 'uvTexture' and the raw
-- 'Texture' constructor still exist and still type-check; scenes that use them are rejected by 'raytrace' with an
-- explicit error (RT_ERR_USER_DEFINED) — the GPU backend has no CPU fallback.
module Graphics.Ray.Texture
  ( Texture(Texture)
  , constantTexture, solidTexture, uvTexture, imageTexture, checkerTexture, noiseTexture, marbleTexture
  , TextureRep(..), textureRep
  ) where

import Graphics.Ray.Core (Color, Point3, Vec3)
import Linear (V2)
import qualified Data.Massiv.Array as A

-- | What the backend sees.
data TextureRep
  = TexConstant Color                               -- ^ Texture.hs:18-19
  | TexChecker Int Int Color Color                  -- ^ Texture.hs:45-53
  | TexImage (A.Matrix A.S Color)                   -- ^ Texture.hs:31-41 (linear RGB, row 0 = top)
  | TexNoise Int Double Vec3 Color Color            -- ^ layers, frequency, shift, c0, c1  (Texture.hs:56-67)
  | TexMarble Vec3 Double Vec3                      -- ^ direction, frequency, shift       (Texture.hs:70-79)
  | TexUser (Point3 -> V2 Double -> Color)          -- ^ arbitrary closure: rejected at commit time

-- | Same constructor name as the reference.  Pattern-matching on it still yields a function for the library's own
-- textures? No: it wraps the closure form; use the smart constructors below for everything the GPU can render.
data Texture = Texture (Point3 -> V2 Double -> Color) | Reified TextureRep

textureRep :: Texture -> TextureRep
textureRep (Reified r) = r
textureRep (Texture f) = TexUser f

constantTexture :: Color -> Texture
constantTexture = Reified . TexConstant

solidTexture :: (Point3 -> Color) -> Texture
solidTexture f = Texture (\p _ -> f p)

uvTexture :: (V2 Double -> Color) -> Texture
uvTexture f = Texture (\_ uv -> f uv)

imageTexture :: A.Matrix A.S Color -> Texture
imageTexture = Reified . TexImage

checkerTexture :: Int -> Int -> Color -> Color -> Texture
checkerTexture nu nv c0 c1 = Reified (TexChecker nu nv c0 c1)

noiseTexture :: Int -> Double -> Vec3 -> Color -> Color -> Texture
noiseTexture k freq shift c0 c1 = Reified (TexNoise k freq shift c0 c1)

marbleTexture :: Vec3 -> Double -> Vec3 -> Texture
marbleTexture dir freq shift = Reified (TexMarble dir freq shift)
