x = 1
if x then y = 2 end
